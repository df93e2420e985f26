"""Byte layout of a serialised PolynomialBatch (the circuit artefact stores the constants +
sigmas commitment and the prover reloads it instead of recommitting).

Restates plonky2 util/serialization/mod.rs `Write::write_polynomial_batch` / `write_merkle_tree` and their `Read`
counterparts from the public crate; like the rest of the oracle this layout is unpinned (no reference bytes exist
in the mount).  usize -> u64 LE, field element -> canonical u64 LE, hash -> 4 elements, bool -> one byte:

    n_polys, then per polynomial { len, coeffs[len] }
    merkle tree: n_leaves, per leaf { len, elements[len] }, n_digests, digests[n_digests][4], cap_height, cap[2^h][4]
    degree_log, rate_bits, blinding

Pure host-side packing (numpy); the device parts come from / go to `bsx_batch_export` / `bsx_batch_import`.
"""
import numpy as np

SALT_SIZE = 4


def _u(x):
    return np.array([x], dtype="<u8").tobytes()


def write_polynomial_batch(coeffs, leaves, digests, cap, degree_log, rate_bits, blinding):
    coeffs, leaves = np.asarray(coeffs, dtype=np.uint64), np.asarray(leaves, dtype=np.uint64)
    digests, cap = np.asarray(digests, dtype=np.uint64).reshape(-1, 4), np.asarray(cap, dtype=np.uint64).reshape(-1, 4)
    out = [_u(coeffs.shape[0])]
    for c in coeffs:
        out += [_u(c.shape[0]), c.astype("<u8").tobytes()]
    rows = np.empty((leaves.shape[0], leaves.shape[1] + 1), dtype="<u8")
    rows[:, 0] = leaves.shape[1]
    rows[:, 1:] = leaves
    cap_height = int(cap.shape[0]).bit_length() - 1
    out += [_u(leaves.shape[0]), rows.tobytes(), _u(digests.shape[0]), digests.astype("<u8").tobytes(), _u(cap_height),
            cap.astype("<u8").tobytes(), _u(degree_log), _u(rate_bits), bytes([1 if blinding else 0])]
    return b"".join(out)


def read_polynomial_batch(data):
    """-> dict(coeffs, leaves, digests, cap, degree_log, rate_bits, blinding); ValueError on malformed input."""
    buf = memoryview(data).cast("B")
    pos = 0

    def take(n_words):
        nonlocal pos
        end = pos + 8 * n_words
        if n_words < 0 or end > len(buf):
            raise ValueError("polynomial batch bytes truncated")
        a = np.frombuffer(buf[pos:end], dtype="<u8")
        pos = end
        return a

    n_polys = int(take(1)[0])
    if n_polys == 0 or n_polys > 1 << 16:
        raise ValueError("polynomial batch bytes: bad polynomial count")
    coeffs, n = [], None
    for _ in range(n_polys):
        ln = int(take(1)[0])
        if n is not None and ln != n:
            raise ValueError("polynomials of different lengths")
        n = ln
        coeffs.append(take(ln))
    n_leaves = int(take(1)[0])
    if n_leaves == 0:
        raise ValueError("polynomial batch bytes: no leaves")
    width = int(take(1)[0])
    pos -= 8
    rows = take(n_leaves * (width + 1)).reshape(n_leaves, width + 1)
    if not (rows[:, 0] == width).all():
        raise ValueError("leaves of different lengths")
    n_dig = int(take(1)[0])
    digests = take(4 * n_dig).reshape(n_dig, 4)
    cap_height = int(take(1)[0])
    if cap_height > 32:
        raise ValueError("polynomial batch bytes: bad cap height")
    cap = take(4 << cap_height).reshape(1 << cap_height, 4)
    degree_log, rate_bits = int(take(1)[0]), int(take(1)[0])
    if pos + 1 != len(buf):
        raise ValueError("polynomial batch bytes: wrong length")
    blinding = buf[pos] != 0
    if degree_log > 40 or rate_bits > 8 or n != 1 << degree_log or n_leaves != n << rate_bits or \
            width != n_polys + (SALT_SIZE if blinding else 0) or n_dig != 2 * (n_leaves - (1 << cap_height)):
        raise ValueError("polynomial batch bytes: inconsistent shape")
    return {"coeffs": np.stack(coeffs).astype(np.uint64), "leaves": np.ascontiguousarray(rows[:, 1:]).astype(np.uint64),
            "digests": digests.astype(np.uint64), "cap": cap.astype(np.uint64), "degree_log": degree_log,
            "rate_bits": rate_bits, "blinding": blinding}


# ---------------------------------------------------------------------------------------------
# FriProof: the flat word stream of bsx_fri_prove_openings <-> plonky2's struct and byte layout
def fri_proof_struct(words, leaf_widths, degree_bits, rate_bits, cap_height, arity_bits, num_query_rounds):
    """Cuts the flat proof (include/bsx_b200.h: layout of bsx_fri_proof_words) into plonky2's `FriProof` fields:
    {commit_phase_merkle_caps: [cap (2^h, 4)], query_round_proofs: [{initial_trees_proof: [(leaf, siblings)],
    steps: [(evals (2^arity, 2), siblings)]}], final_poly: (len, 2), pow_witness}.
    leaf_widths: per oracle, polynomials + salt columns."""
    w = np.asarray(words, dtype=np.uint64).reshape(-1)
    pos = 0

    def take(n, shape=None):
        nonlocal pos
        if pos + n > w.size:
            raise ValueError("FRI proof words truncated")
        a = w[pos:pos + n]
        pos += n
        return a.reshape(shape) if shape else a

    log_nb = degree_bits + rate_bits
    caps = [take(4 << cap_height, (1 << cap_height, 4)) for _ in arity_bits]
    rounds = []
    for _ in range(num_query_rounds):
        init = [(take(width), take(4 * (log_nb - cap_height), (log_nb - cap_height, 4))) for width in leaf_widths]
        steps, lg = [], log_nb
        for ab in arity_bits:
            lg -= ab
            steps.append((take(2 << ab, (1 << ab, 2)), take(4 * (lg - cap_height), (lg - cap_height, 4))))
        rounds.append({"initial_trees_proof": init, "steps": steps})
    final_poly = take(2 << (degree_bits - sum(arity_bits)), (1 << (degree_bits - sum(arity_bits)), 2))
    pow_witness = int(take(1)[0])
    if pos != w.size:
        raise ValueError("FRI proof words: %d trailing words" % (w.size - pos))
    return {"commit_phase_merkle_caps": caps, "query_round_proofs": rounds, "final_poly": final_poly,
            "pow_witness": pow_witness}


def write_fri_proof(proof):
    """plonky2 util/serialization `Write::write_fri_proof` (restated from the public crate, unpinned): caps as
    hashes, per query round the initial-tree openings (`write_field_vec` leaf, `write_merkle_proof` = u8 sibling
    count + hashes) then per step the evals (`write_field_ext_vec`) and its Merkle proof, the final polynomial's
    coefficients, the proof-of-work witness.  Vectors carry no length prefix: the reader knows them from
    `CommonCircuitData` (`read_fri_proof`)."""
    out = []

    def merkle_proof(sib):
        if sib.shape[0] > 255:
            raise ValueError("Merkle proof longer than a u8 length")
        out.append(bytes([sib.shape[0]]))
        out.append(sib.astype("<u8").tobytes())

    for cap in proof["commit_phase_merkle_caps"]:
        out.append(np.asarray(cap).astype("<u8").tobytes())
    for qr in proof["query_round_proofs"]:
        for leaf, sib in qr["initial_trees_proof"]:
            out.append(np.asarray(leaf).astype("<u8").tobytes())
            merkle_proof(np.asarray(sib))
        for evals, sib in qr["steps"]:
            out.append(np.asarray(evals).astype("<u8").tobytes())
            merkle_proof(np.asarray(sib))
    out.append(np.asarray(proof["final_poly"]).astype("<u8").tobytes())
    out.append(np.array([proof["pow_witness"]], dtype="<u8").tobytes())
    return b"".join(out)


def read_fri_proof(data, leaf_widths, degree_bits, rate_bits, cap_height, arity_bits, num_query_rounds):
    """`Read::read_fri_proof`: bytes -> the flat word stream (inverse of write_fri_proof . fri_proof_struct)."""
    buf = memoryview(data).cast("B")
    pos, words = 0, []

    def take(n_words):
        nonlocal pos
        end = pos + 8 * n_words
        if end > len(buf):
            raise ValueError("FRI proof bytes truncated")
        words.append(np.frombuffer(buf[pos:end], dtype="<u8"))
        pos = end

    def merkle_proof(n_sib):
        nonlocal pos
        if pos >= len(buf) or buf[pos] != n_sib:
            raise ValueError("FRI proof bytes: unexpected Merkle proof length")
        pos += 1
        take(4 * n_sib)

    log_nb = degree_bits + rate_bits
    for _ in arity_bits:
        take(4 << cap_height)
    for _ in range(num_query_rounds):
        for width in leaf_widths:
            take(width)
            merkle_proof(log_nb - cap_height)
        lg = log_nb
        for ab in arity_bits:
            lg -= ab
            take(2 << ab)
            merkle_proof(lg - cap_height)
    take(2 << (degree_bits - sum(arity_bits)))
    take(1)
    if pos != len(buf):
        raise ValueError("FRI proof bytes: wrong length")
    return np.concatenate(words).astype(np.uint64)


def splice_initial_openings(fri_words, openings, degree_bits, rate_bits, cap_height, arity_bits, num_query_rounds):
    """The flat proof of a prove over row-sharded oracles: `fri_words` is the stream of
    bsx_fri_proof_from_final_poly with n_oracles = 0 (caps, per query round the FRI steps only, final polynomial,
    witness); `openings` = per oracle (rows (num_query_rounds, width), siblings (num_query_rounds, n_sib, 4)) as the
    owning ranks served them.  Returns the stream bsx_fri_prove_openings produces on one GPU."""
    w = np.asarray(fri_words, dtype=np.uint64).reshape(-1)
    log_nb = degree_bits + rate_bits
    caps_words = len(arity_bits) * (4 << cap_height)
    per_q, lg = 0, log_nb
    for ab in arity_bits:
        lg -= ab
        per_q += (2 << ab) + 4 * (lg - cap_height)
    tail = (2 << (degree_bits - sum(arity_bits))) + 1
    if w.size != caps_words + per_q * num_query_rounds + tail:
        raise ValueError("FRI-only proof stream has the wrong length")
    out = [w[:caps_words]]
    for q in range(num_query_rounds):
        for rows, sib in openings:
            rows, sib = np.asarray(rows, dtype=np.uint64), np.asarray(sib, dtype=np.uint64)
            if sib.shape[1:] != (log_nb - cap_height, 4):
                raise ValueError("initial-tree proof of the wrong length")
            out += [rows[q].reshape(-1), sib[q].reshape(-1)]
        out.append(w[caps_words + q * per_q:caps_words + (q + 1) * per_q])
    out.append(w[w.size - tail:])
    return np.concatenate(out)
