// api_comm.cu -- the sharded commit and prove_openings for ONE process driving G GPUs (SURVEY.md section 8b:
// `bsx_comm_create(n_ranks, devices[], &comm)`): what a Rust prover links against when it wants one
// PolynomialBatch spread over the GPUs of a box.  plonky2 has no multi-device path; the partitioning is the one
// api_shard.cu describes (columns -> ranks for the transforms, leaf rows -> ranks for hashing, the last NTT pass
// stores every word into its owner's row slab over NVLink peer access).
//
// One host thread per rank runs that rank's calls on its own bsx_ctx (contexts are independent; the library has no
// global mutable state), so the G devices work concurrently; the joins between the phases are the barriers the
// exchange needs (all slabs allocated -> all peer stores complete -> hash).  `devices` may name one device several
// times: the ranks then share it (same words, used by the 1-GPU tests), and no kernel ever waits on another rank's
// kernel, so that is safe.  No collective library is involved: the data path is peer stores, the cap and the
// openings are a few KB read back by the host.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>

#include "api_internal.h"

struct bsx_comm {
    int n, log_n_ranks;
    std::vector<bsx_ctx *> ctx;
    std::vector<int> device;
    char err[640];
};

struct bsx_sharded_batch {
    bsx_comm *comm;
    int log_n, n_cols, rate_bits, cap_height;
    int local_cap_height, reduce_levels, chunk_log;  // chunk_log = log2(leaves per rank)
    std::vector<bsx_shard *> shard;
    std::vector<bsx_tree *> tree;
    std::vector<u64> rank_caps;  // n ranks x 2^local_cap_height x 4
    std::vector<u64> cap;        // 2^cap_height x 4
};

static int comm_fail(bsx_comm *c, int code, const char *fmt, ...) {
    if (c) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(c->err, sizeof(c->err), fmt, ap);
        va_end(ap);
    }
    return code;
}

// fn(rank) on one host thread per rank; returns the first failing rank's code and copies its message
static int run_ranks(bsx_comm *c, const std::function<int(int)> &fn) {
    std::vector<int> rc((size_t)c->n, BSX_OK);
    if (c->n == 1) {
        rc[0] = fn(0);
    } else {
        std::vector<std::thread> th;
        for (int r = 0; r < c->n; r++) th.emplace_back([&, r] { rc[r] = fn(r); });
        for (auto &t : th) t.join();
    }
    for (int r = 0; r < c->n; r++)
        if (rc[r] != BSX_OK) {
            snprintf(c->err, sizeof(c->err), "rank %d: %s", r, bsx_last_error(c->ctx[r]));
            return rc[r];
        }
    return BSX_OK;
}

extern "C" int32_t bsx_comm_create(int32_t n_ranks, const int32_t *devices, bsx_comm **out) {
    if (!out || !devices) return BSX_ERR_INVALID_ARG;
    *out = nullptr;
    int lg = 0;
    while ((1 << lg) < n_ranks) lg++;
    if (n_ranks < 1 || n_ranks > BSX_MAX_RANKS || (1 << lg) != n_ranks) return BSX_ERR_INVALID_ARG;
    bsx_comm *c = new bsx_comm();
    c->n = n_ranks;
    c->log_n_ranks = lg;
    c->err[0] = 0;
    for (int r = 0; r < n_ranks; r++) {
        bsx_ctx *x = nullptr;
        int rc = bsx_ctx_create(devices[r], &x);
        if (rc != BSX_OK) {
            for (auto *p : c->ctx) bsx_ctx_destroy(p);
            delete c;
            return rc;
        }
        c->ctx.push_back(x);
        c->device.push_back(devices[r]);
    }
    *out = c;
    return BSX_OK;
}
extern "C" void bsx_comm_destroy(bsx_comm *c) {
    if (!c) return;
    for (auto *p : c->ctx) bsx_ctx_destroy(p);
    delete c;
}
extern "C" int32_t bsx_comm_size(const bsx_comm *c) { return c ? c->n : 0; }
extern "C" bsx_ctx *bsx_comm_ctx(bsx_comm *c, int32_t rank) { return (c && rank >= 0 && rank < c->n) ? c->ctx[rank] : nullptr; }
extern "C" const char *bsx_comm_last_error(const bsx_comm *c) { return c ? c->err : "null communicator"; }

extern "C" void bsx_sharded_free(bsx_sharded_batch *b) {
    if (!b) return;
    for (auto *t : b->tree)
        if (t) bsx_merkle_free(t);
    for (auto *s : b->shard)
        if (s) bsx_shard_free(s);
    delete b;
}

extern "C" int32_t bsx_sharded_from_values_parts(bsx_comm *c, const uint64_t *const *parts, int32_t mem, int32_t log_n,
                                                 int32_t n_cols, int32_t rate_bits, int32_t cap_height,
                                                 int32_t input_is_coeffs, bsx_sharded_batch **out, bsx_timing *timing) {
    if (!c || !out || !parts) return BSX_ERR_INVALID_ARG;
    *out = nullptr;
    bsx_sharded_batch *b = new bsx_sharded_batch();
    b->comm = c;
    b->log_n = log_n;
    b->n_cols = n_cols;
    b->rate_bits = rate_bits;
    b->cap_height = cap_height;
    b->shard.assign((size_t)c->n, nullptr);
    b->tree.assign((size_t)c->n, nullptr);
    std::vector<bsx_timing> tm((size_t)c->n);
    // phase 1: every rank allocates its row slab
    int rc = run_ranks(c, [&](int r) { return bsx_shard_create(c->ctx[r], log_n, n_cols, rate_bits, cap_height, r, c->n, &b->shard[r]); });
    // phase 2: peer pointers, then the transforms; their last pass stores into the owners' slabs
    if (rc == BSX_OK)
        rc = run_ranks(c, [&](int r) {
            for (int p = 0; p < c->n; p++)
                if (p != r) TRY(bsx_shard_set_peer_ptr(b->shard[r], p, bsx_shard_slab_device(b->shard[p])));
            return (int)bsx_shard_transform(b->shard[r], parts[r], mem, input_is_coeffs, 0, timing ? &tm[r] : nullptr);
        });
    // phase 3 (every rank's stores are complete: bsx_shard_transform returns after its stream has drained):
    // leaf hashing and the local cap subtrees
    int32_t lc = 0;
    if (rc == BSX_OK) {
        uint64_t chunk = 0;
        bsx_shard_shape(b->shard[0], nullptr, nullptr, &chunk, &lc);
        b->local_cap_height = lc;
        b->chunk_log = log_n + rate_bits - c->log_n_ranks;
        b->reduce_levels = cap_height < c->log_n_ranks ? c->log_n_ranks - cap_height : 0;
        b->rank_caps.assign((size_t)c->n * ((size_t)4 << lc), 0);
        std::vector<cudaEvent_t> e0((size_t)c->n, nullptr), e1((size_t)c->n, nullptr);
        rc = run_ranks(c, [&](int r) {
            bsx_ctx *x = c->ctx[r];
            if (timing) {
                cudaSetDevice(x->device);
                cudaEventCreate(&e0[r]);
                cudaEventCreate(&e1[r]);
                cudaEventRecord(e0[r], x->stream);
            }
            TRY(bsx_shard_commit(b->shard[r], &b->tree[r]));
            if (timing) {
                cudaEventRecord(e1[r], x->stream);
                cudaEventSynchronize(e1[r]);
                cudaEventElapsedTime(&tm[r].leaf_hash_ms, e0[r], e1[r]);  // leaf hashing + subtrees of this rank
                cudaEventDestroy(e0[r]);
                cudaEventDestroy(e1[r]);
            }
            return (int)bsx_merkle_cap(b->tree[r], (uint64_t *)b->rank_caps.data() + (size_t)r * ((size_t)4 << lc));
        });
    }
    // the cap of the whole batch: local caps in rank order, rank roots hashed up when there are more ranks than entries
    if (rc == BSX_OK) {
        b->cap.assign((size_t)4 << cap_height, 0);
        if (b->reduce_levels) {
            rc = bsx_merkle_reduce(c->ctx[0], (const uint64_t *)b->rank_caps.data(), c->n, b->reduce_levels, (uint64_t *)b->cap.data());
            if (rc != BSX_OK) comm_fail(c, rc, "rank 0: %s", bsx_last_error(c->ctx[0]));
        } else {
            b->cap = b->rank_caps;
        }
    }
    if (rc != BSX_OK) {
        bsx_sharded_free(b);
        return rc;
    }
    if (timing) {  // the slowest rank of every phase
        memset(timing, 0, sizeof(*timing));
        for (int r = 0; r < c->n; r++) {
            timing->ifft_ms = tm[r].ifft_ms > timing->ifft_ms ? tm[r].ifft_ms : timing->ifft_ms;
            timing->lde_ms = tm[r].lde_ms > timing->lde_ms ? tm[r].lde_ms : timing->lde_ms;
            timing->leaf_hash_ms = tm[r].leaf_hash_ms > timing->leaf_hash_ms ? tm[r].leaf_hash_ms : timing->leaf_hash_ms;
        }
        timing->total_ms = timing->ifft_ms + timing->lde_ms + timing->leaf_hash_ms;
    }
    *out = b;
    return BSX_OK;
}

extern "C" int32_t bsx_sharded_from_values(bsx_comm *c, const uint64_t *values_host, int32_t log_n, int32_t n_cols,
                                           int32_t rate_bits, int32_t cap_height, int32_t input_is_coeffs,
                                           bsx_sharded_batch **out, bsx_timing *timing) {
    if (!c || !out || (!values_host && n_cols > 0)) return BSX_ERR_INVALID_ARG;
    if (log_n < 0 || log_n > 24 || n_cols < 0) return comm_fail(c, BSX_ERR_INVALID_ARG, "bad shape");
    // column-major with column stride 2^log_n: rank r's columns [c0, c1) are one contiguous range
    std::vector<const uint64_t *> parts((size_t)c->n);
    for (int r = 0; r < c->n; r++) {
        int32_t c0, c1;
        bsx_shard_column_range(n_cols, r, c->n, &c0, &c1);
        parts[r] = values_host + ((size_t)c0 << log_n);
    }
    return bsx_sharded_from_values_parts(c, parts.data(), BSX_HOST, log_n, n_cols, rate_bits, cap_height, input_is_coeffs,
                                         out, timing);
}

extern "C" int32_t bsx_sharded_cap(const bsx_sharded_batch *b, uint64_t *out_host) {
    if (!b || !out_host) return BSX_ERR_INVALID_ARG;
    memcpy(out_host, b->cap.data(), b->cap.size() * sizeof(u64));
    return BSX_OK;
}
extern "C" bsx_shard *bsx_sharded_shard(bsx_sharded_batch *b, int32_t rank) {
    return (b && rank >= 0 && rank < b->comm->n) ? b->shard[rank] : nullptr;
}
extern "C" bsx_tree *bsx_sharded_tree(bsx_sharded_batch *b, int32_t rank) {
    return (b && rank >= 0 && rank < b->comm->n) ? b->tree[rank] : nullptr;
}
extern "C" int32_t bsx_sharded_proof_len(const bsx_sharded_batch *b) {
    return b ? b->log_n + b->rate_bits - b->cap_height : 0;
}

// MerkleTree::get for global leaf indices: each row comes from the rank that holds it
extern "C" int32_t bsx_sharded_get(bsx_sharded_batch *b, const uint64_t *leaf_idx_host, int32_t k, uint64_t *out_host) {
    if (!b || (k > 0 && (!leaf_idx_host || !out_host))) return BSX_ERR_INVALID_ARG;
    bsx_comm *c = b->comm;
    for (int i = 0; i < k; i++) {
        const uint64_t l = leaf_idx_host[i];
        if (l >> (b->log_n + b->rate_bits)) return comm_fail(c, BSX_ERR_INVALID_ARG, "leaf %llu out of range", (unsigned long long)l);
        const int owner = (int)(l >> b->chunk_log);
        const uint64_t local = l & (((uint64_t)1 << b->chunk_log) - 1);
        int rc = bsx_merkle_get(b->tree[owner], &local, 1, out_host + (size_t)i * b->n_cols);
        if (rc != BSX_OK) return comm_fail(c, rc, "rank %d: %s", owner, bsx_last_error(c->ctx[owner]));
    }
    return BSX_OK;
}

// MerkleTree::prove against the cap of the WHOLE batch: the owner's local path, then (more ranks than cap entries)
// the siblings above its root
extern "C" int32_t bsx_sharded_prove(bsx_sharded_batch *b, uint64_t leaf, uint64_t *siblings_host) {
    if (!b || !siblings_host) return BSX_ERR_INVALID_ARG;
    bsx_comm *c = b->comm;
    if (leaf >> (b->log_n + b->rate_bits)) return comm_fail(c, BSX_ERR_INVALID_ARG, "leaf %llu out of range", (unsigned long long)leaf);
    const int owner = (int)(leaf >> b->chunk_log);
    const uint64_t local = leaf & (((uint64_t)1 << b->chunk_log) - 1);
    int rc = bsx_merkle_prove(b->tree[owner], local, siblings_host);
    if (rc != BSX_OK) return comm_fail(c, rc, "rank %d: %s", owner, bsx_last_error(c->ctx[owner]));
    if (b->reduce_levels) {
        u64 *sib = (u64 *)siblings_host + 4 * (size_t)(b->chunk_log - b->local_cap_height);
        std::vector<u64> level = b->rank_caps, next;
        int idx = owner;
        for (int lv = 0; lv < b->reduce_levels; lv++) {
            memcpy(sib + 4 * (size_t)lv, level.data() + 4 * (size_t)(idx ^ 1), 4 * sizeof(u64));
            next.assign(level.size() / 2, 0);
            rc = bsx_merkle_reduce(c->ctx[0], (const uint64_t *)level.data(), (int)(level.size() / 4), 1, (uint64_t *)next.data());
            if (rc != BSX_OK) return comm_fail(c, rc, "rank 0: %s", bsx_last_error(c->ctx[0]));
            level.swap(next);
            idx >>= 1;
        }
    }
    return BSX_OK;
}

// PolynomialBatch::prove_openings over row-sharded oracles (the call sequence documented in the header, in one call).
extern "C" int32_t bsx_sharded_prove_openings(bsx_comm *c, bsx_sharded_batch *const *oracles, int32_t n_oracles,
                                              const bsx_fri_instance *instance, bsx_challenger *challenger,
                                              const bsx_fri_params *params, uint64_t *proof_host) {
    if (!c || !oracles || n_oracles < 1 || n_oracles > BSX_MAX_ORACLES || !instance || !challenger || !params || !proof_host)
        return BSX_ERR_INVALID_ARG;
    for (int o = 0; o < n_oracles; o++)
        if (!oracles[o] || oracles[o]->comm != c || oracles[o]->log_n != params->degree_bits ||
            oracles[o]->rate_bits != params->rate_bits || oracles[o]->cap_height != params->cap_height)
            return comm_fail(c, BSX_ERR_INVALID_ARG, "oracle %d does not match the FRI parameters", o);
    bsx_ctx *x0 = c->ctx[0];
    uint64_t alpha[2];
    int rc = bsx_challenger_get(x0, challenger, alpha, 2);  // challenger.get_extension_challenge()
    if (rc != BSX_OK) return comm_fail(c, rc, "rank 0: %s", bsx_last_error(x0));
    // every rank: its part of the final polynomial from the coefficient columns it holds
    const size_t n = (size_t)1 << params->degree_bits;
    std::vector<u64> parts((size_t)c->n * 2 * n);
    rc = run_ranks(c, [&](int r) {
        std::vector<bsx_shard *> sh((size_t)n_oracles);
        for (int o = 0; o < n_oracles; o++) sh[o] = oracles[o]->shard[r];
        return (int)bsx_shard_final_poly_part(sh.data(), n_oracles, instance, alpha, (uint64_t *)parts.data() + (size_t)r * 2 * n, BSX_HOST);
    });
    if (rc != BSX_OK) return rc;
    std::vector<u64> fin(2 * n);
    rc = bsx_field_sum(x0, (const uint64_t *)parts.data(), BSX_HOST, c->n, 2 * n, (uint64_t *)fin.data(), BSX_HOST);
    if (rc != BSX_OK) return comm_fail(c, rc, "rank 0: %s", bsx_last_error(x0));
    // commit phase, proof of work, FRI-tree openings (not sharded: < 2 % of the work, serialised by Fiat-Shamir)
    bsx_fri_layout L0, L;
    std::vector<int32_t> widths((size_t)n_oracles);
    for (int o = 0; o < n_oracles; o++) widths[o] = oracles[o]->n_cols;
    if ((rc = bsx_fri_proof_layout(params, nullptr, 0, &L0)) || (rc = bsx_fri_proof_layout(params, widths.data(), n_oracles, &L)))
        return comm_fail(c, rc, "bad FRI parameters");
    std::vector<u64> fri(L0.total_words), idx((size_t)params->num_query_rounds);
    rc = bsx_fri_proof_from_final_poly(x0, nullptr, 0, (const uint64_t *)fin.data(), BSX_HOST, challenger, params, (uint64_t *)fri.data(), (uint64_t *)idx.data());
    if (rc != BSX_OK) return comm_fail(c, rc, "rank 0: %s", bsx_last_error(x0));
    // splice: caps, per query round the initial-tree openings served by the owners, then the FRI steps
    u64 *w = (u64 *)proof_host;
    memcpy(w + L.caps_offset, fri.data() + L0.caps_offset, L.cap_words * (size_t)L.n_rounds * sizeof(u64));
    for (int q = 0; q < params->num_query_rounds; q++) {
        u64 *qd = w + L.query_offset + (size_t)q * L.query_stride;
        for (int o = 0; o < n_oracles; o++) {
            if ((rc = bsx_sharded_get(oracles[o], (const uint64_t *)&idx[q], 1, (uint64_t *)(qd + L.initial_leaf_offset[o])))) return rc;
            if ((rc = bsx_sharded_prove(oracles[o], idx[q], (uint64_t *)(qd + L.initial_proof_offset[o])))) return rc;
        }
        if (L.n_rounds > 0)
            memcpy(qd + L.step_evals_offset[0], fri.data() + L0.query_offset + (size_t)q * L0.query_stride,
                   L0.query_stride * sizeof(u64));
    }
    memcpy(w + L.final_poly_offset, fri.data() + L0.final_poly_offset, (L.final_poly_words + 1) * sizeof(u64));
    return BSX_OK;
}
