// api_proof.cu -- typed view and byte serialisation of the proof objects (host code only; no kernels).
// Replaces plonky2 util/serialization/mod.rs `Write::{write_fri_proof, write_fri_query_round, write_fri_initial_proof,
// write_fri_query_step, write_merkle_proof, write_merkle_cap, write_opening_set, write_proof,
// write_proof_with_public_inputs}` (= `ProofWithPublicInputs::to_bytes`, plonk/proof.rs) for
// F = Goldilocks, D = 2, Poseidon hashes.  usize/field element -> u64 little-endian; a hash is four elements; a
// Merkle proof is one length byte + its siblings; vectors whose length the reader knows from CommonCircuitData carry
// no prefix; the public inputs fill the rest of the buffer (`read_proof_with_public_inputs` reads
// `remaining() / 8` elements).  The layout is restated from the public crate: unpinned, like the oracle.
#include <cstring>

#include "../../include/bsx_b200.h"

static int layout_of(const bsx_fri_params *p, const int32_t *leaf_widths, int n_oracles, bsx_fri_layout *L) {
    if (!p || !L || n_oracles < 0 || n_oracles > BSX_MAX_ORACLES || (n_oracles > 0 && !leaf_widths)) return BSX_ERR_INVALID_ARG;
    if (p->n_arity < 0 || p->n_arity > 16 || p->cap_height < 0 || p->num_query_rounds < 0 || p->degree_bits < 0 ||
        p->rate_bits < 0)
        return BSX_ERR_INVALID_ARG;
    memset(L, 0, sizeof(*L));
    const int log_nb = p->degree_bits + p->rate_bits;
    if (p->cap_height > log_nb) return BSX_ERR_INVALID_ARG;
    L->n_oracles = n_oracles;
    L->n_rounds = p->n_arity;
    L->n_query_rounds = p->num_query_rounds;
    L->cap_words = (uint64_t)4 << p->cap_height;
    L->caps_offset = 0;
    uint64_t pos = L->cap_words * (uint64_t)p->n_arity;
    L->query_offset = pos;
    uint64_t q = 0;
    L->initial_siblings = log_nb - p->cap_height;
    for (int o = 0; o < n_oracles; o++) {
        if (leaf_widths[o] < 0) return BSX_ERR_INVALID_ARG;
        L->initial_leaf_offset[o] = q;
        L->initial_leaf_words[o] = (uint64_t)leaf_widths[o];
        q += (uint64_t)leaf_widths[o];
        L->initial_proof_offset[o] = q;
        q += 4 * (uint64_t)L->initial_siblings;
    }
    int lg = log_nb, sum = 0;
    for (int r = 0; r < p->n_arity; r++) {
        const int ab = p->arity_bits[r];
        if (ab < 1 || ab > 8) return BSX_ERR_INVALID_ARG;
        lg -= ab;
        sum += ab;
        if (lg < p->cap_height) return BSX_ERR_INVALID_ARG;
        L->step_evals_offset[r] = q;
        L->step_evals_words[r] = (uint64_t)2 << ab;
        q += (uint64_t)2 << ab;
        L->step_proof_offset[r] = q;
        L->step_siblings[r] = lg - p->cap_height;
        q += 4 * (uint64_t)(lg - p->cap_height);
    }
    if (sum > p->degree_bits) return BSX_ERR_INVALID_ARG;
    L->query_stride = q;
    pos += q * (uint64_t)p->num_query_rounds;
    L->final_poly_offset = pos;
    L->final_poly_words = (uint64_t)2 << (p->degree_bits - sum);
    pos += L->final_poly_words;
    L->pow_witness_offset = pos;
    L->total_words = pos + 1;
    return BSX_OK;
}

extern "C" int32_t bsx_fri_proof_layout(const bsx_fri_params *params, const int32_t *leaf_widths, int32_t n_oracles,
                                        bsx_fri_layout *out) {
    return layout_of(params, leaf_widths, n_oracles, out);
}

namespace {
struct Writer {  // counts when buf == nullptr
    uint8_t *buf;
    size_t cap, pos;
    bool overflow;
    void bytes(const void *p, size_t n) {
        if (buf) {
            if (pos + n > cap) overflow = true;
            else memcpy(buf + pos, p, n);  // the ABI is little-endian u64 on both sides
        }
        pos += n;
    }
    void words(const uint64_t *w, uint64_t n) { bytes(w, (size_t)n * 8); }
    void u8(uint8_t v) { bytes(&v, 1); }
    void merkle_proof(const uint64_t *sib, int n_sib) {  // write_merkle_proof: u8 length, then the hashes
        u8((uint8_t)n_sib);
        words(sib, 4 * (uint64_t)n_sib);
    }
};

int write_fri(Writer &w, const uint64_t *s, const bsx_fri_layout &L) {
    for (int r = 0; r < L.n_rounds; r++) w.words(s + L.caps_offset + (uint64_t)r * L.cap_words, L.cap_words);
    if (L.initial_siblings > 255) return BSX_ERR_UNSUPPORTED;  // upstream: try_into::<u8>().expect(..)
    for (int qi = 0; qi < L.n_query_rounds; qi++) {
        const uint64_t *q = s + L.query_offset + (uint64_t)qi * L.query_stride;
        for (int o = 0; o < L.n_oracles; o++) {
            w.words(q + L.initial_leaf_offset[o], L.initial_leaf_words[o]);  // write_field_vec(leaf)
            w.merkle_proof(q + L.initial_proof_offset[o], L.initial_siblings);
        }
        for (int r = 0; r < L.n_rounds; r++) {
            w.words(q + L.step_evals_offset[r], L.step_evals_words[r]);  // write_field_ext_vec(evals)
            w.merkle_proof(q + L.step_proof_offset[r], L.step_siblings[r]);
        }
    }
    w.words(s + L.final_poly_offset, L.final_poly_words);
    w.words(s + L.pow_witness_offset, 1);
    return BSX_OK;
}
}  // namespace

extern "C" int32_t bsx_fri_proof_bytes(const uint64_t *proof_words, const bsx_fri_params *params,
                                       const int32_t *leaf_widths, int32_t n_oracles, uint8_t *out, size_t out_cap,
                                       size_t *out_len) {
    bsx_fri_layout L;
    int rc = layout_of(params, leaf_widths, n_oracles, &L);
    if (rc) return rc;
    if (!proof_words || !out_len) return BSX_ERR_INVALID_ARG;
    Writer w = {out, out_cap, 0, false};
    if ((rc = write_fri(w, proof_words, L))) return rc;
    *out_len = w.pos;
    return w.overflow ? BSX_ERR_INVALID_ARG : BSX_OK;
}

extern "C" int32_t bsx_proof_with_public_inputs_bytes(const bsx_proof_parts *parts, const bsx_fri_params *params,
                                                      const int32_t *leaf_widths, int32_t n_oracles, uint8_t *out,
                                                      size_t out_cap, size_t *out_len) {
    bsx_fri_layout L;
    int rc = layout_of(params, leaf_widths, n_oracles, &L);
    if (rc) return rc;
    if (!parts || !out_len || !parts->wires_cap || !parts->plonk_zs_partial_products_cap || !parts->quotient_polys_cap ||
        !parts->opening_proof_words || (parts->n_public_inputs > 0 && !parts->public_inputs))
        return BSX_ERR_INVALID_ARG;
    Writer w = {out, out_cap, 0, false};
    // write_proof: wires_cap, plonk_zs_partial_products_cap, quotient_polys_cap, openings, opening_proof
    w.words(parts->wires_cap, L.cap_words);
    w.words(parts->plonk_zs_partial_products_cap, L.cap_words);
    w.words(parts->quotient_polys_cap, L.cap_words);
    // write_opening_set: constants, plonk_sigmas, wires, plonk_zs, plonk_zs_next, partial_products, quotient_polys
    // (lookup_zs / lookup_zs_next of later revisions sit after plonk_zs_next and are empty without lookup tables)
    for (int k = 0; k < BSX_OPENING_SET_FIELDS; k++) {
        if (parts->n_openings[k] < 0 || (parts->n_openings[k] > 0 && !parts->openings[k])) return BSX_ERR_INVALID_ARG;
        w.words(parts->openings[k], 2 * (uint64_t)parts->n_openings[k]);
    }
    if ((rc = write_fri(w, parts->opening_proof_words, L))) return rc;
    if (parts->n_public_inputs < 0) return BSX_ERR_INVALID_ARG;
    w.words(parts->public_inputs, (uint64_t)parts->n_public_inputs);  // write_field_vec(public_inputs)
    *out_len = w.pos;
    return w.overflow ? BSX_ERR_INVALID_ARG : BSX_OK;
}
