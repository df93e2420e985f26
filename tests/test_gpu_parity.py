"""Tier 2: every CUDA kernel bit-exact against the CPU oracle, through the C ABI.  Integer field
arithmetic: comparisons are exact equality, no tolerances."""
import numpy as np
import pytest

from conftest import P, splitmix_field
import __graft_entry__ as G

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    return G.load_package()


@pytest.fixture(scope="module")
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()


def _add_mod(a, b):
    """(a + b) mod p on canonical uint64 arrays without leaving uint64."""
    with np.errstate(over="ignore"):
        t = a + b
        wrap = (t < a) | (t >= np.uint64(P))
        return np.where(wrap, t - np.uint64(P), t)


EDGE = [0, 1, 2, P - 1, P - 2, 2**32 - 1, 2**32, 2**32 + 1, 2**63, P - 2**32, 0xFFFFFFFF00000000, 0xFFFFFFFF]


def test_poseidon_kat_on_device(ctx):
    z = ctx.poseidon_permute(np.zeros(12, dtype=np.uint64))
    assert "%016x" % int(z[0]) == "3c18a9786cb0b359" and "%016x" % int(z[11]) == "1792b1c4342109d7"
    i = ctx.poseidon_permute(np.arange(12, dtype=np.uint64))
    assert "%016x" % int(i[0]) == "d64e1e3efc5b8e9e"
    m = ctx.poseidon_permute(np.full(12, P - 1, dtype=np.uint64))
    assert "%016x" % int(m[0]) == "be0085cfc57a8357"


def test_poseidon_random_and_adversarial(ctx, oracle):
    st = splitmix_field(42, (4096, 12))
    # rows built from reduction edge values: products hit the borrow/carry paths of reduce128
    adv = np.array([[EDGE[(i + j) % len(EDGE)] for j in range(12)] for i in range(64)], dtype=np.uint64)
    st[:64] = adv
    got = ctx.poseidon_permute(st)
    want = oracle.poseidon(st)
    assert (got == want).all()


# 13: rows of 2^6 words (ntt_pass_lo); 14..19: rows of 2^7..2^9 words, the tile of pass LO moved by TMA (ntt_pass_lo_tma)
@pytest.mark.parametrize("log_n", [0, 1, 2, 5, 10, 12, 13, 14, 15, 16, 17, 19])
def test_ntt_kinds(ctx, oracle, log_n):
    n_cols = 3
    v = splitmix_field(100 + log_n, (n_cols, 1 << log_n))
    if log_n >= 2:
        v[0, :4] = np.array([0, P - 1, 2**32 - 1, 2**32], dtype=np.uint64)
    got = ctx.ntt(v, "fft")
    assert all((got[c] == oracle.fft(v[c])).all() for c in range(n_cols))
    got = ctx.ntt(v, "ifft")
    assert all((got[c] == oracle.ifft(v[c])).all() for c in range(n_cols))
    got = ctx.ntt(v, "coset_fft", shift=7)
    assert all((got[c] == oracle.coset_fft(v[c], 7)).all() for c in range(n_cols))
    got = ctx.ntt(v, "coset_fft", shift=pow(7, 16, P))
    assert (got[1] == oracle.coset_fft(v[1], pow(7, 16, P))).all()


@pytest.mark.parametrize("log_n,n_cols", [(13, 9), (16, 8), (14, 17), (18, 9)])
def test_ntt_many_columns_full_factor_table(ctx, oracle, log_n, n_cols):
    """>= 8 columns at two-pass sizes take the cached full post-factor table path of pass HI."""
    v = splitmix_field(300 + log_n, (n_cols, 1 << log_n))
    for kind, ref in (("fft", oracle.fft), ("ifft", oracle.ifft), ("coset_fft", lambda a: oracle.coset_fft(a, 7))):
        got = ctx.ntt(v, kind)
        for c in (0, n_cols // 2, n_cols - 1):
            assert (got[c] == ref(v[c])).all(), (kind, c)


@pytest.mark.parametrize("log_n,n_cols", [(20, 8), (23, 2), (24, 1)])
def test_ntt_roundtrip_up_to_max_size(ctx, log_n, n_cols):
    """ifft(fft(x)) == x and coset_fft linearity at sizes the oracle would take minutes for."""
    v = splitmix_field(900 + log_n, (n_cols, 1 << log_n))
    assert (ctx.ntt(ctx.ntt(v, "fft"), "ifft") == v).all()
    w = splitmix_field(901 + log_n, (1, 1 << log_n))
    a, b, c = ctx.ntt(v[:1], "coset_fft"), ctx.ntt(w, "coset_fft"), ctx.ntt(_add_mod(v[:1], w), "coset_fft")
    assert (_add_mod(a, b) == c).all()


def test_ntt_roundtrip_large(ctx):
    v = splitmix_field(7, (2, 1 << 20))
    assert (ctx.ntt(ctx.ntt(v, "fft"), "ifft") == v).all()


@pytest.mark.parametrize("n_leaves,leaf_len,cap_height", [
    (1, 7, 0), (2, 3, 0), (2, 3, 1), (16, 4, 2), (64, 5, 0), (64, 135, 4), (1024, 8, 4), (1024, 9, 10),
    (2048, 32, 4), (4096, 20, 0), (1 << 14, 17, 4),
])
def test_merkle_tree_vs_oracle(pkg, ctx, oracle, n_leaves, leaf_len, cap_height):
    leaves = splitmix_field(n_leaves + leaf_len, (n_leaves, leaf_len))
    t = pkg.MerkleTree(ctx, leaves, cap_height)
    digests, cap = oracle.merkle_build(leaves, cap_height)
    assert (t.cap == cap).all()
    log_leaves = n_leaves.bit_length() - 1
    for i in sorted({0, n_leaves - 1, n_leaves // 3}):
        assert (t.get(i) == leaves[i]).all()
        sib = t.prove(i)
        assert (sib == oracle.merkle_prove(digests, log_leaves, cap_height, i)).all()
        assert oracle.merkle_verify(leaves[i], i, sib, cap)
    want_p2 = oracle.merkle_digests_to_plonky2(digests, log_leaves, cap_height)
    assert (t.digests.reshape(-1) == want_p2.reshape(-1)).all()
    t.close()


def test_merkle_errors_match_upstream_panics(pkg, ctx):
    leaves = splitmix_field(1, (8, 5))
    with pytest.raises(pkg.BsxError, match="cap_height=4 should be at most"):
        pkg.MerkleTree(ctx, leaves, 4)
    with pytest.raises(pkg.BsxError, match="power of two"):
        pkg.MerkleTree(ctx, leaves[:6], 0)
    t = pkg.MerkleTree(ctx, leaves, 1)
    with pytest.raises(pkg.BsxError, match="out of range"):
        t.prove(8)


@pytest.mark.parametrize("log_n,n_cols,rate_bits,cap_height", [
    (0, 1, 1, 0), (3, 2, 1, 0), (5, 7, 2, 1), (8, 135, 3, 4), (10, 9, 3, 4), (12, 20, 3, 4), (13, 16, 1, 4),
    (14, 5, 4, 0), (11, 3, 0, 2),
])
def test_batch_from_values_vs_oracle(pkg, ctx, oracle, log_n, n_cols, rate_bits, cap_height):
    vals = splitmix_field(0x5EED0000 + log_n, (n_cols, 1 << log_n))
    b = pkg.PolynomialBatch.from_values(ctx, vals, rate_bits, False, cap_height)
    want = oracle.batch_from_values(vals, rate_bits, cap_height)
    assert (b.cap == want["cap"]).all()
    ex = b.export()
    assert (ex["coeffs"] == want["coeffs"]).all()
    assert (ex["leaves"] == want["leaves"]).all()
    log_nb = log_n + rate_bits
    if want["digests"] is not None:
        assert (ex["digests"].reshape(-1) ==
                oracle.merkle_digests_to_plonky2(want["digests"], log_nb, cap_height).reshape(-1)).all()
    # get_lde_values(index, step) = leaves[reverse_bits(index * step)]
    for idx, step in ((0, 1), ((1 << log_n) - 1, 1 << rate_bits), (1, 1)):
        if idx * step >> log_nb:
            continue
        leaf = int(oracle.lib().orc_reverse_bits(idx * step, log_nb))
        assert (b.get_lde_values(idx, step) == want["leaves"][leaf]).all()
    b.close()


def test_config1_standard_recursion_commit(pkg, ctx, oracle):
    """BASELINE.json configs[0]: 2^14 x 135, rate_bits 3, cap_height 4."""
    vals = splitmix_field(0x5EED0001, (135, 1 << 14))
    b = pkg.PolynomialBatch.from_values(ctx, vals, 3, False, 4, timing=True)
    want = oracle.batch_from_values(vals, 3, 4, want=("digests",))
    assert (b.cap == want["cap"]).all()
    tree = b.merkle_tree
    for i in (0, 12345, (1 << 17) - 1):
        assert (tree.prove(i) == oracle.merkle_prove(want["digests"], 17, 4, i)).all()
    assert b.timing.total_ms > 0
    b.close()


def test_edge_inputs_zero_and_minus_one(pkg, ctx, oracle):
    for fill in (0, P - 1):
        vals = np.full((6, 1 << 9), fill, dtype=np.uint64)
        b = pkg.PolynomialBatch.from_values(ctx, vals, 3, False, 4)
        assert (b.cap == oracle.batch_from_values(vals, 3, 4, want=())["cap"]).all()
        b.close()


def test_from_coeffs_and_salt(pkg, ctx, oracle):
    log_n, n_cols, rate_bits = 9, 6, 3
    coeffs = splitmix_field(77, (n_cols, 1 << log_n))
    b = pkg.PolynomialBatch.from_coeffs(ctx, coeffs, rate_bits, False, 4)
    want = oracle.batch_from_values(coeffs, rate_bits, 4, from_coeffs=True)
    assert (b.cap == want["cap"]).all()
    assert (b.polynomials == coeffs).all()
    b.close()
    salt = splitmix_field(78, (4, 1 << (log_n + rate_bits)))
    b = pkg.PolynomialBatch.from_values(ctx, coeffs, rate_bits, True, 4, salt=salt)
    want = oracle.batch_from_values(coeffs, rate_bits, 4, salt=salt)
    assert (b.cap == want["cap"]).all()
    assert b.shape()["n_salt"] == 4
    assert (b.export()["leaves"] == want["leaves"]).all()
    # salt is stripped by get_lde_values
    assert b.get_lde_values(3, 1).shape == (n_cols,)
    b.close()
    # device-drawn salt: caps differ from the unsalted commit, tree is still self-consistent
    b = pkg.PolynomialBatch.from_values(ctx, coeffs, rate_bits, True, 4)
    ex = b.export()
    _, cap = oracle.merkle_build(ex["leaves"], 4)
    assert (cap == b.cap).all() and ex["leaves"].shape[1] == n_cols + 4
    assert (ex["leaves"] < np.uint64(P)).all()
    b.close()


def test_linearity_full_size(pkg, ctx):
    """Size-independent property at a size the oracle would take minutes for: LDE is linear, so
    lde(a + b) == lde(a) + lde(b) row by row (checked on sampled rows)."""
    log_n, n_cols = 18, 16
    a = splitmix_field(1, (n_cols, 1 << log_n))
    b = splitmix_field(2, (n_cols, 1 << log_n))
    s = ((a.astype(object) + b.astype(object)) % P).astype(np.uint64)
    ba, bb, bs = (pkg.PolynomialBatch.from_values(ctx, x, 3, False, 4) for x in (a, b, s))
    for idx in (0, 1, 99999, (1 << 21) - 1):
        ra, rb, rs = (t.get_lde_values(idx) for t in (ba, bb, bs))
        assert (((ra.astype(object) + rb.astype(object)) % P).astype(np.uint64) == rs).all()
    for t in (ba, bb, bs):
        t.close()


@pytest.mark.parametrize("blinding", [False, True])
def test_batch_import_and_bytes_roundtrip(pkg, ctx, oracle, blinding):
    """export -> bsx_batch_import gives a batch that answers every accessor identically; the serialised bytes
    equal the bytes written from the oracle's parts."""
    ser = __import__(pkg.__name__ + ".serialization", fromlist=["x"])
    log_n, n_cols, rate_bits, cap_height = 9, 7, 2, 3
    vals = splitmix_field(401, (n_cols, 1 << log_n))
    salt = splitmix_field(402, (4, 1 << (log_n + rate_bits))) if blinding else None
    a = pkg.PolynomialBatch.from_values(ctx, vals, rate_bits, blinding, cap_height, salt=salt)
    ob = oracle.batch_from_values(vals, rate_bits, cap_height, salt=salt)
    data = a.to_bytes()
    want = ser.write_polynomial_batch(ob["coeffs"], ob["leaves"],
                                      oracle.merkle_digests_to_plonky2(ob["digests"], log_n + rate_bits, cap_height),
                                      ob["cap"], log_n, rate_bits, blinding)
    assert data == want
    b = pkg.PolynomialBatch.from_bytes(ctx, data)
    assert a.shape() == b.shape()
    assert (a.cap == b.cap).all() and (a.polynomials == b.polynomials).all()
    ea, eb = a.export(), b.export()
    for k in ("coeffs", "leaves", "digests", "cap"):
        assert (ea[k] == eb[k]).all(), k
    for i in (0, 1, 777, (1 << (log_n + rate_bits)) - 1):
        assert (a.get_lde_values(i) == b.get_lde_values(i)).all()
        pa, pb = a.merkle_tree.prove(i), b.merkle_tree.prove(i)
        assert (np.asarray(pa) == np.asarray(pb)).all()
        assert oracle.merkle_verify(b.merkle_tree.get(i), i, pb, b.cap)
    z = splitmix_field(403, (2,))
    assert (a.eval_commitment(z) == b.eval_commitment(z)).all()
    with pytest.raises(pkg.BsxError):
        pkg.PolynomialBatch.from_bytes(ctx, data[:-9])
    a.close()
    b.close()


def test_fri_proof_from_imported_batch(pkg, ctx, oracle):
    """prove_openings on a restored batch (coefficients held in natural order) = on the freshly built one."""
    log_n, rate_bits, cap_height = 10, 1, 1
    vals = [splitmix_field(411 + i, (w, 1 << log_n)) for i, w in enumerate((4, 3))]
    fresh = [pkg.PolynomialBatch.from_values(ctx, v, rate_bits, False, cap_height) for v in vals]
    restored = [pkg.PolynomialBatch.from_bytes(ctx, f.to_bytes()) for f in fresh]
    zeta = splitmix_field(420, (1, 2))
    polys = [[(o, j) for o, v in enumerate(vals) for j in range(v.shape[0])]]
    inst = pkg.FriInstanceInfo(zeta, polys)
    params = pkg.FriConfig(rate_bits=rate_bits, cap_height=cap_height, proof_of_work_bits=6,
                           num_query_rounds=5).fri_params(log_n)
    proofs = []
    for bs in (fresh, restored):
        ch = pkg.Challenger(ctx)
        ch.observe_elements(splitmix_field(421, (8,)))
        proofs.append(pkg.prove_openings(inst, bs, ch, params))
    assert (np.asarray(proofs[0]) == np.asarray(proofs[1])).all()
    for x in fresh + restored:
        x.close()


def test_transpose_tall_matrix(pkg, ctx, oracle):
    """MerkleTree::new on 2^22 row-major leaves: the transpose grid is one-dimensional (2^17 row tiles)."""
    n, w = 1 << 22, 2
    leaves = splitmix_field(431, (n, w))
    t = pkg.MerkleTree(ctx, leaves, 0)
    for i in (0, 12345, n - 1):
        assert (t.get(i) == leaves[i]).all()
        assert oracle.merkle_verify(leaves[i], i, t.prove(i), t.cap)
    t.close()


def test_get_lde_values_packed(pkg, ctx):
    log_n, rate_bits = 10, 3
    vals = splitmix_field(501, (6, 1 << log_n))
    b = pkg.PolynomialBatch.from_values(ctx, vals, rate_bits, True, 2)
    for start, step, width in ((0, 1, 4), (12, 8, 8), ((1 << (log_n + rate_bits)) - 4, 1, 4), (5, 3, 1)):
        got = b.get_lde_values_packed(start, step, width)
        assert got.shape == (6, width)
        for i in range(width):
            assert (got[:, i] == b.get_lde_values(start + i, step)).all()
    with pytest.raises(pkg.BsxError):
        b.get_lde_values_packed((1 << (log_n + rate_bits)) - 2, 1, 4)
    b.close()


def test_noncanonical_inputs_are_residues(pkg, ctx, oracle):
    """plonky2's GoldilocksField(u64) may hold a non-canonical representative (x + p < 2^64 for x < 2^32 - 1); a
    `Vec<F>` handed over by the Rust shim can contain such words.  Every entry point treats a word as its residue."""
    log_n, n_cols, rate_bits, cap_height = 9, 6, 2, 3
    vals = splitmix_field(601, (n_cols, 1 << log_n))
    small = (vals % np.uint64(2**32 - 1))
    vals[:, ::3] = small[:, ::3]                      # canonical small values ...
    alias = vals.copy()
    alias[:, ::3] = small[:, ::3] + np.uint64(P)      # ... and their non-canonical aliases
    assert (alias[:, ::3] >= np.uint64(P)).all()
    for mk in (pkg.PolynomialBatch.from_values, pkg.PolynomialBatch.from_coeffs):
        a, b = mk(ctx, vals, rate_bits, False, cap_height), mk(ctx, alias, rate_bits, False, cap_height)
        assert (a.cap == b.cap).all()
        ea, eb = a.export(), b.export()
        assert (ea["leaves"] == eb["leaves"]).all() and (ea["digests"] == eb["digests"]).all()
        assert (eb["coeffs"] % np.uint64(P) == ea["coeffs"]).all()
        a.close()
        b.close()
    rows_a, rows_b = vals.T.copy(), alias.T.copy()
    ta, tb = pkg.MerkleTree(ctx, rows_a, 2), pkg.MerkleTree(ctx, rows_b, 2)
    assert (ta.cap == tb.cap).all()
    ta.close()
    tb.close()
    st = np.stack([small[0, :12], small[0, :12] + np.uint64(P)])
    out = ctx.poseidon_permute(st)
    assert (out[0] == out[1]).all()
    assert (ctx.ntt(alias[:2], "fft") == ctx.ntt(vals[:2], "fft")).all()


@pytest.mark.parametrize("log_n,n_cols", [(13, 40), (14, 17), (9, 5), (13, 16)])
def test_separately_allocated_pageable_columns(pkg, ctx, oracle, log_n, n_cols):
    """`Vec<PolynomialValues<F>>`: one pageable heap vector per column (pointer-array entry point, staged through the
    pinned ring by host threads) and a pageable contiguous buffer give the cap of the pinned/contiguous path."""
    vals = splitmix_field(701 + log_n, (n_cols, 1 << log_n))
    want = oracle.batch_from_values(vals, 2, 3, want=())["cap"]
    vecs = [vals[c].copy() for c in range(n_cols)]
    for mk, src in ((pkg.PolynomialBatch.from_values, vecs), (pkg.PolynomialBatch.from_values, vals)):
        for rep in range(2):  # second pass reuses the ring
            b = mk(ctx, src, 2, False, 3, timing=True)
            assert (b.cap == want).all()
            b.close()
    co = pkg.PolynomialBatch.from_coeffs(ctx, vecs, 2, False, 3)
    assert (co.cap == oracle.batch_from_values(vals, 2, 3, want=(), from_coeffs=True)["cap"]).all()
    assert (co.polynomials == vals).all()
    co.close()
    with pytest.raises(pkg.BsxError):
        pkg.PolynomialBatch.from_values(ctx, [vals[0], vals[1][:8]], 2, False, 3)


def test_export_leaves_in_slabs(pkg, ctx, oracle):
    """Row-major leaf export runs in ~32 MiB row slabs through the pinned ring: five slabs at 2^17 x 135, so the
    ring (four slots) wraps and the two device buffers are reused."""
    import torch
    vals = splitmix_field(801, (135, 1 << 14))
    b = pkg.PolynomialBatch.from_values(ctx, vals, 3, False, 4)
    want = oracle.batch_from_values(vals, 3, 4, want=("leaves",))
    ex = b.export(coeffs=False, digests=False)
    assert (ex["leaves"] == want["leaves"]).all()
    # pinned destination: DMA straight into the caller's buffer
    pinned = torch.empty((1 << 17, 135), dtype=torch.int64).pin_memory()
    out = pinned.numpy().view(np.uint64)
    ctx.check(pkg.lib().bsx_batch_export(b._h, None, out.ctypes.data, None))
    assert (out == want["leaves"]).all()
    b.close()
