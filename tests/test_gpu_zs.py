"""Zs and partial products of the permutation argument (plonk/prover.rs `all_wires_permutation_partial_products`)
through the C ABI against the oracle, and into the commitment that follows."""
import numpy as np
import pytest

from conftest import P, splitmix_field
from test_oracle import _perm_instance
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


@pytest.mark.parametrize("log_n,n_routed,deg,swaps", [
    (0, 3, 2, 0), (3, 5, 2, 6), (6, 8, 8, 40), (8, 80, 8, 500), (10, 80, 8, 3000), (11, 33, 7, 900), (13, 12, 3, 5000)])
def test_partial_products_and_zs_vs_oracle(ctx, oracle, log_n, n_routed, deg, swaps):
    wires, sigmas, k_is, _ = _perm_instance(700 + log_n, log_n, n_routed, swaps)
    betas = splitmix_field(1, (2,))
    gammas = splitmix_field(2, (2,))
    got = ctx.partial_products_and_zs(wires, sigmas, k_is, betas, gammas, deg)
    want = oracle.partial_products_and_zs(wires, sigmas, k_is, betas, gammas, deg)
    assert got.shape == want.shape == (2 * -(-n_routed // deg), 1 << log_n)
    assert (got == want).all()
    assert (got[:2, 0] == 1).all()


def test_random_sigmas_and_noncanonical_words(ctx, oracle):
    """No copy constraints behind the sigmas (the product does not close) and input words >= p."""
    n_routed, log_n = 20, 9
    wires = splitmix_field(5, (n_routed, 1 << log_n))
    sigmas = splitmix_field(6, (n_routed, 1 << log_n))
    k_is = splitmix_field(7, (n_routed,))
    got = ctx.partial_products_and_zs(wires, sigmas, k_is, [11, 12, 13], [14, 15, 16], 8)
    assert (got == oracle.partial_products_and_zs(wires, sigmas, k_is, [11, 12, 13], [14, 15, 16], 8)).all()
    big = wires.copy()
    big[big < np.uint64(2**32 - 1)] += np.uint64(P)  # same residues, non-canonical words
    assert (ctx.partial_products_and_zs(big, sigmas, k_is, [11, 12, 13], [14, 15, 16], 8) == got).all()


def test_long_scan_many_blocks(ctx, oracle):
    """2^19 rows: the scan of the block products runs over more than one 256-block pass."""
    n_routed, log_n = 3, 19
    wires = splitmix_field(50, (n_routed, 1 << log_n))
    sigmas = splitmix_field(51, (n_routed, 1 << log_n))
    k_is = splitmix_field(52, (n_routed,))
    got = ctx.partial_products_and_zs(wires, sigmas, k_is, [21, 22], [23, 24], 2)
    want = oracle.partial_products_and_zs(wires, sigmas, k_is, [21, 22], [23, 24], 2)
    assert got.shape == (4, 1 << log_n) and (got == want).all()


def test_device_in_device_out_feeds_the_zs_commitment(pkg, ctx, oracle):
    import torch
    log_n, n_routed, deg = 12, 80, 8
    wires, sigmas, k_is, _ = _perm_instance(41, log_n, n_routed, 4000)
    betas, gammas = splitmix_field(3, (2,)), splitmix_field(4, (2,))
    want = oracle.partial_products_and_zs(wires, sigmas, k_is, betas, gammas, deg)
    dw = torch.from_numpy(wires.view(np.int64)).to("cuda:0")
    ds = torch.from_numpy(sigmas.view(np.int64)).to("cuda:0")
    out = torch.empty((20, 1 << log_n), dtype=torch.int64, device="cuda:0")
    ctx.partial_products_and_zs(None, None, k_is, betas, gammas, deg, out_device_ptr=out.data_ptr(),
                                device_ptrs=(dw.data_ptr(), ds.data_ptr()), shape=(n_routed, 1 << log_n))
    assert (out.cpu().numpy().view(np.uint64) == want).all()
    b = pkg.PolynomialBatch.from_values(ctx, None, 3, False, 4, device_ptr=out.data_ptr(), shape=(20, 1 << log_n))
    assert (b.cap == oracle.batch_from_values(want, 3, 4, want=())["cap"]).all()
    b.close()


def test_errors_match_upstream_panics(pkg, ctx, oracle):
    wires, sigmas, k_is, _ = _perm_instance(12, 4, 9, 0)
    beta, gamma = 7, 9
    wires[2, 5] = np.uint64((-(beta * int(sigmas[2, 5]) + gamma)) % P)
    used = ctx.mem_used()
    with pytest.raises(pkg.BsxError):  # batch_multiplicative_inverse: tried to invert zero
        ctx.partial_products_and_zs(wires, sigmas, k_is, [beta], [gamma], 4)
    with pytest.raises(pkg.BsxError):  # quotient_chunk_products: max_degree > 1
        ctx.partial_products_and_zs(wires, sigmas, k_is, [1], [2], 1)
    assert ctx.mem_used() == used
