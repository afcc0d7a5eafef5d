"""ctb_kin_getNetProductionRates_ddX_noAlloc (include/ctb_clib.h): the allocation-free variant of the
Jacobian tile kernel on the library's static scratch array, against the allocating entry point
ctb_kin_getNetProductionRates_ddX on the same device buffers, two calls in a row (the array keeps the
previous call's contents). Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from conftest import MECHS

pytestmark = pytest.mark.gpu

#: the tolerance of the ddX parity tests (tests/test_gpu_kinetics.py), relative to max |J| of the state;
#: the two entry points run the same per-state arithmetic, so the difference is expected to be 0
TOL = 1e-12


def _states(K, n, seed):
    import torch
    rng = np.random.default_rng(seed)
    T = rng.uniform(800, 2500, n)
    P = 101325.0 * rng.uniform(1, 50, n)
    Y = rng.uniform(0, 1, (n, K))
    Y /= Y.sum(axis=1, keepdims=True)
    return [torch.from_numpy(x).cuda() for x in (T, P, Y)]


def _both(lib, h, K, T, P, Y, jac_a, jac_b):
    """The existing entry point into jac_a, the allocation-free one into jac_b, through the C ABI."""
    import torch
    from cantera_b200 import clib
    n = T.numel()
    st = torch.cuda.current_stream().cuda_stream
    clib.check(lib.ctb_kin_setStream(h, st))
    clib.check(lib.ctb_kin_getNetProductionRates_ddX(h, n, T.data_ptr(), P.data_ptr(), Y.data_ptr(), K,
                                                     jac_a.data_ptr(), jac_a.numel(), clib.DEVICE))
    rc = lib.ctb_kin_getNetProductionRates_ddX_noAlloc(h, n, T.data_ptr(), P.data_ptr(), Y.data_ptr(), K,
                                                       jac_b.data_ptr(), jac_b.numel())
    if rc == -1 and "does not fit the Jacobian tile kernel" in clib.last_error():
        return None  # the documented rejection of mechanisms beyond the tile kernel's shared memory
    clib.check(rc)
    torch.cuda.synchronize()
    a, b = jac_a[:n].cpu().numpy(), jac_b[:n].cpu().numpy()
    scale = np.abs(a).max(axis=(1, 2), keepdims=True)
    assert np.all(np.isfinite(b))
    return float(np.max(np.abs(a - b) / scale))


@pytest.mark.parametrize("mech,n1,n2", [("gri30", 20000, 1237), ("h2o2", 50000, 31), ("nDodecane", 3000, 5000)])
def test_no_alloc_matches_allocating_entry_point(mech, n1, n2):
    import torch
    from cantera_b200 import BatchedGasKinetics, clib
    kin = BatchedGasKinetics(MECHS[mech], device=0)
    lib, h, K = clib.lib(), kin._h, kin.n_species
    nmax = max(n1, n2)
    jac_a = torch.empty((nmax, K, K), dtype=torch.float64, device="cuda")
    jac_b = torch.full((nmax, K, K), float("nan"), dtype=torch.float64, device="cuda")
    # first call: more tiles than the array holds at once for gri30 / nDodecane (capped grid, strided tiles)
    T, P, Y = _states(K, n1, 11)
    err1 = _both(lib, h, K, T, P, Y, jac_a, jac_b)
    if err1 is None:
        assert mech == "nDodecane"  # the config mechanisms gri30 and h2o2 must be served
        kin.close()
        return
    assert err1 <= TOL, "first call: %g" % err1
    # second call in a row, other states and another batch size: the scratch holds the first call's rows
    jac_b.fill_(float("nan"))
    T, P, Y = _states(K, n2, 12)
    err2 = _both(lib, h, K, T, P, Y, jac_a, jac_b)
    assert err2 <= TOL, "second call: %g" % err2
    # the Python mirror takes the same route
    out = torch.empty((n2, K, K), dtype=torch.float64, device="cuda")
    kin.net_production_rates_ddX_no_alloc(T, P, Y, out)
    torch.cuda.synchronize()
    assert torch.equal(out, jac_b[:n2])
    kin.close()


def test_no_alloc_makes_no_device_allocation():
    """Free device memory does not move across calls (cudaMemGetInfo before / after, other sizes in between)."""
    import torch
    from cantera_b200 import BatchedGasKinetics, clib
    kin = BatchedGasKinetics(MECHS["gri30"], device=0)
    lib, h, K = clib.lib(), kin._h, kin.n_species
    n = 4096
    T, P, Y = _states(K, n, 5)
    jac = torch.empty((n, K, K), dtype=torch.float64, device="cuda")

    def call(m):
        clib.check(lib.ctb_kin_setStream(h, torch.cuda.current_stream().cuda_stream))
        clib.check(lib.ctb_kin_getNetProductionRates_ddX_noAlloc(h, m, T.data_ptr(), P.data_ptr(), Y.data_ptr(), K,
                                                               jac.data_ptr(), jac.numel()))
        torch.cuda.synchronize()
    call(n)  # set-up: resolves the array's address, loads the kernel
    free0 = torch.cuda.mem_get_info()[0]
    for m in (n, 17, 1, 2048, n):
        call(m)
    assert torch.cuda.mem_get_info()[0] == free0
    kin.close()


def test_no_alloc_rejects_beyond_the_maximum():
    from cantera_b200 import BatchedGasKinetics, clib
    kin = BatchedGasKinetics(MECHS["h2o2"], device=0)
    lib, h, K = clib.lib(), kin._h, kin.n_species
    n = clib.NOALLOC_MAX_STATES + 1
    # rejected before any pointer is touched
    rc = lib.ctb_kin_getNetProductionRates_ddX_noAlloc(h, n, C.c_void_p(8), C.c_void_p(8), C.c_void_p(8), K,
                                                       C.c_void_p(8), n * K * K)
    assert rc == -1
    with pytest.raises(clib.CtbError, match="exceed the maximum of %d" % clib.NOALLOC_MAX_STATES):
        clib.check(rc)
    # the array size check of the other getters
    rc = lib.ctb_kin_getNetProductionRates_ddX_noAlloc(h, 4, C.c_void_p(8), C.c_void_p(8), C.c_void_p(8), K,
                                                       C.c_void_p(8), 4 * K * K - 1)
    assert rc == -1
    with pytest.raises(clib.CtbError, match="too small"):
        clib.check(rc)
    kin.close()
