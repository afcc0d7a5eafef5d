"""SolutionArray-shaped front end and the library-level multi-GPU split (cantera_b200/solution_array.py).
Two handles on the same device stand in for two GPUs where the box has only one: the slicing, the
host threads and the permuted reactor shares are the same code."""
import numpy as np
import pytest

from conftest import MECHS, golden

pytestmark = pytest.mark.gpu


def _devices():
    import torch
    n = torch.cuda.device_count()
    return [0, 1] if n >= 2 else [0, 0]


def test_solution_array_shapes_and_values():
    from cantera_b200 import BatchedGasKinetics, BatchedSolutionArray
    states = BatchedSolutionArray(MECHS["gri30"], shape=(7, 9), devices=[0])
    K, R = states.n_species, states.n_reactions
    rng = np.random.default_rng(2)
    T = rng.uniform(900, 2400, (7, 9))
    Y = rng.uniform(0, 1, (7, 9, K))
    Y /= Y.sum(axis=-1, keepdims=True)
    states.TPY = T, 5 * 101325.0, Y           # scalar pressure broadcasts like SolutionArray's setter
    assert states.T.shape == (7, 9) and states.P.shape == (7, 9) and states.Y.shape == (7, 9, K)
    w = states.net_production_rates
    assert w.shape == (7, 9, K) and states.forward_rates_of_progress.shape == (7, 9, R)
    assert states.net_production_rates_ddX.shape == (7, 9, K, K)
    kin = BatchedGasKinetics(MECHS["gri30"])
    ref = kin.net_production_rates(T.ravel(), np.full(63, 5 * 101325.0), Y.reshape(63, K))
    assert np.array_equal(w.reshape(63, K), ref)
    kin.close()
    states.close()


def test_multi_gpu_kinetics_matches_single():
    from cantera_b200 import BatchedGasKinetics, MultiGpuKinetics
    multi = MultiGpuKinetics(MECHS["gri30"], devices=_devices())
    kin = BatchedGasKinetics(MECHS["gri30"])
    K = kin.n_species
    rng = np.random.default_rng(4)
    n = 10007                                   # not a multiple of the tile or of the device count
    T = rng.uniform(800, 2500, n)
    P = 101325.0 * rng.uniform(1, 50, n)
    Y = rng.uniform(0, 1, (n, K))
    Y /= Y.sum(axis=1, keepdims=True)
    assert np.array_equal(multi.net_production_rates(T, P, Y), kin.net_production_rates(T, P, Y))
    assert np.array_equal(multi.net_rates_of_progress(T, P, Y), kin.net_rates_of_progress(T, P, Y))
    assert np.array_equal(multi.net_production_rates_ddX(T[:300], P[:300], Y[:300]), kin.net_production_rates_ddX(T[:300], P[:300], Y[:300]))
    kin.close()
    multi.close()


def test_multi_gpu_reactor_net_matches_single():
    from cantera_b200 import BatchedGasKinetics, BatchReactorNet, MultiGpuKinetics, MultiGpuReactorNet
    g = golden("ignite", "h2o2")
    N, K, R = g["dims"]
    T, P, Y = g["T"], g["P"], g["Y"].reshape(N, K)
    multi = MultiGpuKinetics(MECHS["h2o2"], devices=_devices())
    mnet = MultiGpuReactorNet(multi, N)
    assert sorted(np.concatenate(mnet.shares).tolist()) == list(range(N))
    mnet.set_state(T, P, Y)
    mnet.set_ignition_monitor(400.0)
    mnet.advance(1e-3)
    Tm, Ym = mnet.get_state()
    kin = BatchedGasKinetics(MECHS["h2o2"])
    net = BatchReactorNet(kin, N)
    net.set_state(T, P, Y)
    net.set_ignition_monitor(400.0)
    net.advance(1e-3)
    Ts, Ys = net.get_state()
    # every reactor is integrated independently of its neighbours: same bits wherever it runs
    assert np.array_equal(Tm, Ts) and np.array_equal(Ym, Ys)
    assert np.array_equal(mnet.status(), net.status())
    assert np.array_equal(mnet.ignition_delay(), net.ignition_delay(), equal_nan=True)
    assert mnet.solver_stats().shape == (N, 8)
    net.close(); kin.close(); mnet.close(); multi.close()
