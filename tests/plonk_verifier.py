"""A restatement of plonky2's verifier (plonk/verifier.rs `verify_with_challenges`, plonk/get_challenges.rs) over the
oracle: test infrastructure that ACCEPTS or REJECTS a ProofWithPublicInputs produced by the device prover.  It shares no
code with the prover path: the vanishing polynomial comes from the oracle's extension-field evaluation of the opening
set (orc_vanishing_at_ext), the transcript from the oracle's challenger, FRI from orc_fri_verify."""
import numpy as np

P = 0xFFFFFFFF00000001


def _ext_mul(a, b):
    a0, a1, b0, b1 = int(a[0]), int(a[1]), int(b[0]), int(b[1])
    return ((a0 * b0 + 7 * a1 * b1) % P, (a0 * b1 + a1 * b0) % P)


def verify(oracle, circuit, pw, aux):
    """circuit: the oracle's dict view (CircuitDescription.as_dict()); pw: ProofWithPublicInputs; aux: what the verifier
    holds besides the proof (circuit digest, constants_sigmas cap, FRI parameters, oracle widths).  -> (ok, reason)."""
    pr, fp = pw.proof, aux["fri_params"]
    cfg = fp.config
    nc, n_bits = circuit["num_challenges"], circuit["degree_bits"]
    pih = oracle.hash_no_pad(np.asarray(pw.public_inputs, dtype=np.uint64))
    ch = oracle.Challenger()
    ch.observe(np.asarray(aux["circuit_digest"], dtype=np.uint64))
    ch.observe(pih)
    ch.observe(np.asarray(pr.wires_cap, dtype=np.uint64).reshape(-1))
    betas = [ch.get_challenge() for _ in range(nc)]
    gammas = [ch.get_challenge() for _ in range(nc)]
    ch.observe(np.asarray(pr.plonk_zs_partial_products_cap, dtype=np.uint64).reshape(-1))
    alphas = [ch.get_challenge() for _ in range(nc)]
    ch.observe(np.asarray(pr.quotient_polys_cap, dtype=np.uint64).reshape(-1))
    zeta = ch.get_extension_challenge()
    os_ = pr.openings
    # vanishing(zeta) == Z_H(zeta) * reduce_with_powers(quotient chunk evaluations, zeta^n)
    van = oracle.vanishing_at_ext(circuit, zeta, os_, pih, betas, gammas, alphas)
    zn = (int(zeta[0]), int(zeta[1]))
    for _ in range(n_bits):
        zn = _ext_mul(zn, zn)
    z_h = ((zn[0] - 1) % P, zn[1])
    f = circuit["quotient_degree_factor"]
    q = np.asarray(os_.quotient_polys, dtype=np.uint64).reshape(-1, 2)
    for a in range(nc):
        acc = (0, 0)
        for t in reversed(q[a * f:(a + 1) * f]):
            acc = _ext_mul(acc, zn)
            acc = ((acc[0] + int(t[0])) % P, (acc[1] + int(t[1])) % P)
        if _ext_mul(z_h, acc) != (int(van[a][0]), int(van[a][1])):
            return False, "vanishing polynomial identity fails for challenge %d" % a
    # challenger.observe_openings, then FRI
    fri_openings = os_.to_fri_openings()
    for batch in fri_openings:
        ch.observe(np.asarray(batch, dtype=np.uint64).reshape(-1))
    widths = aux["widths"]
    n = 1 << n_bits
    caps = [aux["constants_sigmas_cap"], pr.wires_cap, pr.plonk_zs_partial_products_cap, pr.quotient_polys_cap]
    shells = [{"coeffs": np.zeros((w, n), dtype=np.uint64), "leaves": np.zeros((1, w), dtype=np.uint64), "digests": None,
               "cap": np.asarray(cap, dtype=np.uint64)} for w, cap in zip(widths, caps)]
    g = pow(1753635133440165772, 1 << (32 - n_bits), P)
    gz = np.array([int(zeta[0]) * g % P, int(zeta[1]) * g % P], dtype=np.uint64)
    polys = [[(o, j) for o, w in enumerate(widths) for j in range(w)], [(2, j) for j in range(nc)]]
    flat = np.concatenate([np.asarray(b, dtype=np.uint64).reshape(-1) for b in fri_openings])
    ok = oracle.fri_verify(shells, np.stack([np.asarray(zeta, dtype=np.uint64), gz]), polys, flat, ch, n_bits, cfg.rate_bits,
                           cfg.cap_height, cfg.proof_of_work_bits, cfg.num_query_rounds, list(fp.reduction_arity_bits),
                           np.asarray(pr.opening_proof, dtype=np.uint64))
    return (True, "ok") if ok else (False, "FRI verification fails")
