"""Mixture-averaged transport (cantera_b200/csrc/transport.cu).
CPU: the numpy restatement oracle/transport_np.py against the reference's own outputs (tests/golden/trans_*:
oracle/ref_build/gen_golden.cpp `trans`), and the product tables against the polynomials in the goldens.
GPU: the kernel through the C ABI against the same goldens: 1e-12 relative for viscosity and conductivity,
1e-12 of the row's largest value for the mixture diffusion coefficients. The diffusion coefficient of a species
whose mole fraction is 1 to rounding is excluded: the reference's (W - x_k W_k) / sum_{j != k} is 0/1e-19-sized
rounding noise there (its own value is exactly 0 or a large number depending on the last bit)."""
import numpy as np
import pytest

from conftest import golden

CASES = ["h2o2", "gri30"]


def _compare(g, mu, lam, dmix, tol):
    N, K, _ = g["dims"]
    Y = g["Y"].reshape(N, K)
    ref = g["mixdiff"].reshape(N, K)
    assert np.max(np.abs(mu / g["viscosity"] - 1)) <= tol
    assert np.max(np.abs(lam / g["conductivity"] - 1)) <= tol
    err = np.abs(dmix - ref) / np.abs(ref).max(axis=1, keepdims=True)
    err[Y > 1 - 1e-9] = 0.0
    assert np.max(err) <= tol


@pytest.mark.parametrize("case", CASES)
def test_restatement_pinned_to_the_reference(case):
    from oracle.transport_np import mixture_transport
    g = golden("trans", case)
    N, K, _ = g["dims"]
    mu, lam, dmix = mixture_transport(g, g["T"], g["P"], g["Y"].reshape(N, K))
    _compare(g, mu, lam, dmix, 1e-13)
    assert np.all(mu > 0) and np.all(lam > 0)


@pytest.mark.parametrize("case", CASES)
def test_product_tables_hold_the_reference_polynomials(case):
    from cantera_b200.ctbm import read_arrays
    from cantera_b200.transport import transport_path
    t = read_arrays(transport_path(case))
    g = golden("trans", case)
    for k in ("mol_weights", "visc_poly", "cond_poly", "diff_poly"):
        assert np.array_equal(t[k], g[k]), k


def test_bridge_extracts_transport_in_process():
    """ctb_trans_new (include/ctb_cantera.h): newSolution(..., "mixture-averaged") by the reference in this process,
    polynomials through its public getters, then ctb_trans_newFromArrays. Without a GPU the call must get as far
    as the device check of the constructor (the extraction itself has then succeeded); with one it yields a
    handle whose results equal those of the table file."""
    import os
    from conftest import ROOT
    from cantera_b200 import clib
    b = clib.bridge()
    ref_dir = os.path.join(ROOT, "oracle", "_ref", "data")
    if b is None or not os.path.exists(os.path.join(ref_dir, "h2o2.yaml")):
        pytest.skip("libctb_cantera.so / the reference's data files are not on this machine")
    clib.check(b.ctb_addDataDirectory(ref_dir.encode()))
    h = b.ctb_trans_new(os.path.join(ref_dir, "h2o2.yaml").encode(), b"ohmech", 0)
    if h == -1:
        assert "no usable CUDA device" in clib.last_error()
        return
    from cantera_b200.transport import BatchedMixTransport
    g = golden("trans", "h2o2")
    N, K, _ = g["dims"]
    assert clib.check(clib.lib().ctb_trans_nSpecies(h)) == K
    clib.check(clib.lib().ctb_trans_del(h))
    tr, tr2 = BatchedMixTransport("h2o2"), BatchedMixTransport.from_yaml(os.path.join(ref_dir, "h2o2.yaml"), "ohmech")
    a = tr.mixture_transport(g["T"], g["P"], g["Y"].reshape(N, K))
    c = tr2.mixture_transport(g["T"], g["P"], g["Y"].reshape(N, K))
    # the fits are recomputed by the reference on THIS host (libm variants may differ from the machine that wrote
    # the table file in the last bits), so not bit for bit
    assert all(np.allclose(x, y, rtol=1e-9, atol=0.0) for x, y in zip(a, c))
    # a phase without the mixture-averaged model is refused with the reference's message
    assert b.ctb_trans_new(os.path.join(ref_dir, "no-such-file.yaml").encode(), b"", 0) == -1


@pytest.mark.gpu
def test_gpu_bridge_transport_matches_table_file():
    test_bridge_extracts_transport_in_process()


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_gpu_mixture_transport(case):
    from cantera_b200.transport import BatchedMixTransport
    g = golden("trans", case)
    N, K, _ = g["dims"]
    Y = g["Y"].reshape(N, K)
    tr = BatchedMixTransport(case)
    assert tr.n_species == K
    mu, lam, dmix = tr.mixture_transport(g["T"], g["P"], Y)
    _compare(g, mu, lam, dmix, 1e-12)
    # any subset of the outputs, ragged leading dimension, device memory
    assert np.array_equal(tr.viscosity(g["T"], g["P"], Y), mu)
    assert np.array_equal(tr.mix_diff_coeffs(g["T"], g["P"], Y), dmix)
    import torch
    Yd = torch.zeros((N, K + 3), dtype=torch.float64, device="cuda")
    Yd[:, :K] = torch.from_numpy(Y).cuda()
    mu_d, lam_d, d_d = tr.mixture_transport(torch.from_numpy(g["T"]).cuda(), torch.from_numpy(g["P"]).cuda(), Yd[:, :K])
    torch.cuda.synchronize()
    assert np.array_equal(mu_d.cpu().numpy(), mu) and np.array_equal(lam_d.cpu().numpy(), lam)
    assert np.array_equal(d_d.cpu().numpy(), dmix)
    # a larger batch: every state independent of its position in the batch
    reps = 37
    mu2 = tr.viscosity(np.tile(g["T"], reps), np.tile(g["P"], reps), np.tile(Y, (reps, 1)))
    assert np.array_equal(mu2.reshape(reps, N), np.tile(mu, (reps, 1)))
    tr.close()


@pytest.mark.gpu
def test_gpu_transport_errors():
    from cantera_b200 import clib
    from cantera_b200.transport import BatchedMixTransport
    tr = BatchedMixTransport("h2o2")
    K = tr.n_species
    rc = clib.lib().ctb_trans_getMixtureTransport(tr._h, 1, 8, 8, 8, K - 1, 8, None, None, K, clib.HOST)
    assert rc == -1 and "ldY is smaller" in clib.last_error()
    rc = clib.lib().ctb_trans_getMixtureTransport(tr._h, 1, None, None, None, K, None, None, None, K, clib.HOST)
    assert rc == -1 and "null state pointer" in clib.last_error()
    tr.close()
    with pytest.raises(clib.CtbError):
        BatchedMixTransport({"mol_weights": np.ones(3), "visc_poly": np.ones(15), "cond_poly": np.ones(15), "diff_poly": np.ones(7)})
