"""Tier 1: the CPU oracle against every anchor available offline."""
import numpy as np
import pytest

from conftest import P, splitmix_field

KAT = {
    "zeros": ([0] * 12, "3c18a9786cb0b359 c4055e3364a246c3 7953db0ab48808f4 c71603f33a1144ca d7709673896996dc 46a84e87642f44ed d032648251ee0b3c 1c687363b207df62 df8565563e8045fe 40f5b37ff4254dae d070f637b431067c 1792b1c4342109d7"),
    "iota": (list(range(12)), "d64e1e3efc5b8e9e 53666633020aaa47 d40285597c6a8825 613a4f81e81231d2 414754bfebd051f0 cb1f8980294a023f 6eb2a9e4d54a9d0f 1902bc3af467e056 f045d5eafdc6021f e4150f77caaa3be5 c9bfd01d39b50cce 5c0a27fcb0e1459b"),
    "neg_one": ([P - 1] * 12, "be0085cfc57a8357 d95af71847d05c09 cf55a13d33c1c953 95803a74f4530e82 fcd99eb30a135df1 e095905e913a3029 de0392461b42919b 7d3260e24e81d031 10d3d0465d9deaa0 a87571083dfc2a47 e18263681e9958f8 e28e96f1ae5e60d3"),
}


def test_field_constants(oracle):
    L = oracle.lib()
    assert P == 2**64 - 2**32 + 1
    assert pow(2, 64, P) == 2**32 - 1 and pow(2, 96, P) == P - 1
    w32 = L.orc_root_of_unity(32)
    assert w32 == 1753635133440165772 == pow(7, (P - 1) >> 32, P)
    assert pow(w32, 1 << 31, P) == P - 1
    assert pow(7, (P - 1) // 2, P) == P - 1  # 7 is a non-residue: x^2 - 7 irreducible
    for k in (1, 5, 20):
        assert L.orc_root_of_unity(k) == pow(w32, 1 << (32 - k), P)


def test_field_ops_vs_python(oracle):
    L = oracle.lib()
    rng = np.random.default_rng(1)
    edge = [0, 1, 2, P - 1, P - 2, 2**32 - 1, 2**32, 2**32 + 1, 2**63, P - 2**32, 0xFFFFFFFF00000000]
    vals = edge + [int(x) for x in rng.integers(0, P, 50, dtype=np.uint64)]
    for a in vals:
        for b in vals[:20]:
            assert L.orc_mul(a, b) == a * b % P
            assert L.orc_add(a, b) == (a + b) % P
            assert L.orc_sub(a, b) == (a - b) % P
    for a in vals[1:12]:
        assert L.orc_mul(a, L.orc_inv(a)) == 1


@pytest.mark.parametrize("name", list(KAT))
def test_poseidon_kat(oracle, name):
    inp, want = KAT[name]
    got = oracle.poseidon(np.array(inp, dtype=np.uint64))
    assert " ".join("%016x" % int(x) for x in got) == want


def test_poseidon_vs_python_naive(oracle):
    import importlib.util, os
    spec = importlib.util.spec_from_file_location(
        "gen", os.path.join(os.path.dirname(__file__), "..", "tools", "gen_poseidon_constants.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    rc = gen.round_constants()
    L = oracle.lib()
    assert [L.orc_round_constants()[i] for i in range(360)] == rc
    st = splitmix_field(7, (4, 12))
    got = oracle.poseidon(st)
    for i in range(4):
        assert [int(x) for x in got[i]] == gen.poseidon_naive([int(x) for x in st[i]], rc)


def test_hashing_modes(oracle):
    x = splitmix_field(3, (135,))
    # hash_or_noop: short inputs are zero-extended, not hashed
    assert list(oracle.hash_or_noop(x[:3])) == [int(x[0]), int(x[1]), int(x[2]), 0]
    assert list(oracle.hash_or_noop(x[:4])) == [int(v) for v in x[:4]]
    # overwrite-mode sponge, no padding: restate by hand with the permutation
    for ln in (5, 8, 9, 16, 135):
        st = np.zeros(12, dtype=np.uint64)
        for off in range(0, ln, 8):
            chunk = x[off:min(off + 8, ln)]
            st[:chunk.size] = chunk
            st = oracle.poseidon(st)
        assert list(oracle.hash_no_pad(x[:ln])) == list(st[:4])
    l, r = x[:4], x[4:8]
    st = np.concatenate([l, r, np.zeros(4, dtype=np.uint64)])
    assert list(oracle.two_to_one(l, r)) == list(oracle.poseidon(st)[:4])


def _eval_poly(coeffs, x):
    acc = 0
    for c in reversed([int(v) for v in coeffs]):
        acc = (acc * x + c) % P
    return acc


@pytest.mark.parametrize("log_n", [0, 1, 3, 6])
def test_fft_vs_naive_evaluation(oracle, log_n):
    n = 1 << log_n
    c = splitmix_field(11 + log_n, (n,))
    w = oracle.lib().orc_root_of_unity(log_n)
    got = oracle.fft(c)
    assert [int(v) for v in got] == [_eval_poly(c, pow(w, i, P)) for i in range(n)]
    assert list(oracle.ifft(got)) == list(c)


@pytest.mark.parametrize("log_n,rate_bits", [(3, 1), (4, 3), (5, 2), (2, 4)])
def test_lde_is_coset_evaluation(oracle, log_n, rate_bits):
    n = 1 << log_n
    c = splitmix_field(21, (n,))
    w = oracle.lib().orc_root_of_unity(log_n + rate_bits)
    got = oracle.lde(c, rate_bits)
    assert [int(v) for v in got] == [_eval_poly(c, 7 * pow(w, i, P) % P) for i in range(n << rate_bits)]


def test_merkle_proofs_verify(oracle):
    leaves = splitmix_field(5, (64, 11))
    for cap_height in (0, 2, 6):
        digests, cap = oracle.merkle_build(leaves, cap_height)
        assert cap.shape == (1 << cap_height, 4)
        for i in (0, 1, 31, 63):
            sib = oracle.merkle_prove(digests, 6, cap_height, i)
            assert sib.shape == (6 - cap_height, 4)
            assert oracle.merkle_verify(leaves[i], i, sib, cap)
            bad = leaves[i].copy()
            bad[0] ^= np.uint64(1)
            assert not oracle.merkle_verify(bad, i, sib, cap)
    # cap_height == log2(leaves): the cap is the leaf digests
    _, cap = oracle.merkle_build(leaves, 6)
    assert list(cap[5]) == list(oracle.hash_or_noop(leaves[5]))
    with pytest.raises(ValueError):
        oracle.merkle_build(leaves, 7)


def test_merkle_root_by_hand(oracle):
    leaves = splitmix_field(6, (4, 3))  # <= 4 elements: noop leaves
    _, cap = oracle.merkle_build(leaves, 0)
    pad = lambda r: np.concatenate([r, np.zeros(1, dtype=np.uint64)])
    l = oracle.two_to_one(pad(leaves[0]), pad(leaves[1]))
    r = oracle.two_to_one(pad(leaves[2]), pad(leaves[3]))
    assert list(cap[0]) == list(oracle.two_to_one(l, r))


def test_plonky2_digest_layout(oracle):
    leaves = splitmix_field(8, (16, 9))
    digests, cap = oracle.merkle_build(leaves, 1)
    p2 = oracle.merkle_digests_to_plonky2(digests, 4, 1).reshape(-1, 4)
    # per cap subtree: 2^(3+1)-2 = 14 digests; the middle pair are the children of the cap node
    assert p2.shape == (28, 4)
    for c in range(2):
        sub = p2[14 * c:14 * (c + 1)]
        assert list(oracle.two_to_one(sub[6], sub[7])) == list(cap[c])
        # leftmost pair of the buffer = first two leaf digests of the subtree
        assert list(sub[0]) == list(oracle.hash_or_noop(leaves[8 * c]))
        assert list(sub[1]) == list(oracle.hash_or_noop(leaves[8 * c + 1]))
        assert list(oracle.two_to_one(sub[0], sub[1])) == list(sub[2])


def test_batch_from_values_composition(oracle):
    log_n, n_cols, rate_bits, cap_height = 5, 7, 2, 1
    vals = splitmix_field(9, (n_cols, 1 << log_n))
    out = oracle.batch_from_values(vals, rate_bits, cap_height)
    nb_log = log_n + rate_bits
    for c in range(n_cols):
        assert list(out["coeffs"][c]) == list(oracle.ifft(vals[c]))
        ldec = oracle.lde(out["coeffs"][c], rate_bits)
        rev = [int(oracle.lib().orc_reverse_bits(l, nb_log)) for l in range(1 << nb_log)]
        assert list(out["leaves"][:, c]) == list(ldec[rev])
    digests, cap = oracle.merkle_build(out["leaves"], cap_height)
    assert (cap == out["cap"]).all() and (digests == out["digests"]).all()
    # from_coeffs agrees with from_values
    out2 = oracle.batch_from_values(out["coeffs"], rate_bits, cap_height, from_coeffs=True)
    assert (out2["cap"] == out["cap"]).all()


def test_batch_salt_columns(oracle):
    log_n, n_cols, rate_bits = 3, 2, 1
    vals = splitmix_field(10, (n_cols, 1 << log_n))
    salt = splitmix_field(12, (4, 1 << (log_n + rate_bits)))
    out = oracle.batch_from_values(vals, rate_bits, 0, salt=salt)
    assert out["leaves"].shape == (16, 6)
    rev = [int(oracle.lib().orc_reverse_bits(l, 4)) for l in range(16)]
    assert (out["leaves"][:, 2:] == salt[:, rev].T).all()


def test_polynomial_batch_bytes_roundtrip(oracle):
    """Host-side wire format (serialization.py) on oracle-produced parts: write -> read is the identity."""
    import __graft_entry__ as G
    pkg = G.load_package()
    ser = __import__(pkg.__name__ + ".serialization", fromlist=["x"])
    vals = splitmix_field(77, (3, 32))
    salt = splitmix_field(78, (4, 64))
    for s in (None, salt):
        ob = oracle.batch_from_values(vals, 1, 2, salt=s)
        dg = oracle.merkle_digests_to_plonky2(ob["digests"], 6, 2)
        data = ser.write_polynomial_batch(ob["coeffs"], ob["leaves"], dg, ob["cap"], 5, 1, s is not None)
        n_lv, w = 64, 3 + (4 if s is not None else 0)
        assert len(data) == 8 * (1 + 3 * 33 + 1 + n_lv * (w + 1) + 1 + 4 * 2 * (64 - 4) + 1 + 16 + 2) + 1
        back = ser.read_polynomial_batch(data)
        assert (back["coeffs"] == ob["coeffs"]).all() and (back["leaves"] == ob["leaves"]).all()
        assert (back["digests"] == dg).all() and (back["cap"] == ob["cap"]).all()
        assert (back["degree_log"], back["rate_bits"], back["blinding"]) == (5, 1, s is not None)
        for bad in (data[:-1], data[:100], data + b"\0", b""):
            with pytest.raises(ValueError):
                ser.read_polynomial_batch(bad)


def test_coset_ifft_inverts_coset_fft_and_interpolates(oracle):
    """PolynomialValues::coset_ifft: the inverse of PolynomialCoeffs::coset_fft, and the interpolant on shift * H."""
    for log_n in (0, 1, 4, 9):
        c = splitmix_field(40 + log_n, (1 << log_n,))
        for shift in (7, 0x123456789ABCDEF):
            assert (oracle.coset_ifft(oracle.coset_fft(c, shift), shift) == c).all()
    n, shift = 8, 7
    c = [int(x) for x in splitmix_field(5, (n,))]
    w = pow(7, (P - 1) // n, P)
    vals = np.array([sum(ci * pow(shift * pow(w, k, P) % P, i, P) for i, ci in enumerate(c)) % P for k in range(n)],
                    dtype=np.uint64)
    assert [int(x) for x in oracle.coset_ifft(vals, shift)] == c


def test_quotient_chunks_trim_and_panic(oracle):
    n, q, factor = 16, 3, 6
    coeffs = np.zeros((2, n << q), dtype=np.uint64)
    coeffs[:, :factor * n] = splitmix_field(8, (2, factor * n))
    vals = np.stack([oracle.coset_fft(c, 7) for c in coeffs])
    ch = oracle.quotient_chunks(vals, 4, factor)
    assert ch.shape == (2 * factor, n) and (ch.reshape(2, -1) == coeffs[:, :factor * n]).all()
    coeffs[0, -1] = 5
    with pytest.raises(ValueError):
        oracle.quotient_chunks(np.stack([oracle.coset_fft(c, 7) for c in coeffs]), 4, factor)


def _perm_instance(seed, log_n, n_routed, swaps):
    """Routed wire values with copy constraints and the sigma values of the matching permutation:
    sigma_j(x_i) = k_{j'} x_{i'} for the position (j', i') that (j, i) is wired to."""
    n = 1 << log_n
    rng = np.random.default_rng(seed)
    wires = splitmix_field(seed, (n_routed, n)).copy()
    k_is = np.array([pow(7, j, P) for j in range(n_routed)], dtype=np.uint64)  # distinct coset representatives
    w = pow(7, (P - 1) >> log_n, P)
    xs = [pow(w, i, P) for i in range(n)]
    target = {(j, i): (j, i) for j in range(n_routed) for i in range(n)}
    for _ in range(swaps):
        a = (int(rng.integers(n_routed)), int(rng.integers(n)))
        b = (int(rng.integers(n_routed)), int(rng.integers(n)))
        if a == b or target[a] != a or target[b] != b:
            continue
        target[a], target[b] = b, a
        wires[b] = wires[a]
    sigmas = np.array([[int(k_is[target[(j, i)][0]]) * xs[target[(j, i)][1]] % P for i in range(n)]
                       for j in range(n_routed)], dtype=np.uint64)
    return wires, sigmas, k_is, xs


def test_partial_products_and_zs_against_direct_formula(oracle):
    """wires_permutation_partial_products_and_zs restated with Python integers on a tiny instance."""
    log_n, n_routed, deg = 3, 5, 2
    wires, sigmas, k_is, xs = _perm_instance(11, log_n, n_routed, 6)
    betas, gammas = [0x1234567, 99], [0xABCDEF01, 5]
    got = oracle.partial_products_and_zs(wires, sigmas, k_is, betas, gammas, deg)
    n, n_chunks = 1 << log_n, 3
    assert got.shape == (2 * n_chunks, n)
    for c, (beta, gamma) in enumerate(zip(betas, gammas)):
        z = 1
        for i in range(n):
            assert int(got[c, i]) == z
            acc = z
            for m in range(n_chunks):
                for j in range(m * deg, min(n_routed, (m + 1) * deg)):
                    num = (int(wires[j, i]) + beta * int(k_is[j]) * xs[i] + gamma) % P
                    den = (int(wires[j, i]) + beta * int(sigmas[j, i]) + gamma) % P
                    acc = acc * num * pow(den, P - 2, P) % P
                if m < n_chunks - 1:
                    assert int(got[2 + c * (n_chunks - 1) + m, i]) == acc
            z = acc
        assert z == 1  # the copy constraints hold, so the grand product closes


def test_partial_products_identity_permutation_and_zero_denominator(oracle):
    wires, sigmas, k_is, xs = _perm_instance(12, 4, 9, 0)
    got = oracle.partial_products_and_zs(wires, sigmas, k_is, [3, 4], [5, 6], 4)
    assert got.shape == (2 * 3, 16) and (got == 1).all()
    beta, gamma = 7, 9
    wires[2, 5] = np.uint64((-(beta * int(sigmas[2, 5]) + gamma)) % P)
    with pytest.raises(ZeroDivisionError):
        oracle.partial_products_and_zs(wires, sigmas, k_is, [beta], [gamma], 4)
