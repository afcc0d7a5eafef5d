"""Host emulation of the device code (tests/emul): the per-state arithmetic of the thread-per-state
kernels (cantera_b200/csrc/rates_thread.cuh, reactor_thread.cuh) and the resumable BDF controller of
the batch integrator (bdf_batch.cuh) are __host__ __device__, so the packer's run table, the
fast/general reaction paths and the controller's state machine are checked here WITHOUT a GPU:
rates against the oracle (1e-12, cancellation-scaled), integrations against the reference's
CVODES golden vectors (same tolerances as tests/test_gpu_reactor.py)."""
import os
import shutil
import struct
import subprocess

import numpy as np
import pytest

from conftest import MECHS, ROOT, golden

EMUL = os.path.join(ROOT, "tests", "emul")
BUILD = os.path.join(EMUL, "_build")
NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"

pytestmark = pytest.mark.skipif(not os.path.exists(NVCC), reason="nvcc not available")


def _build(name):
    os.makedirs(BUILD, exist_ok=True)
    exe = os.path.join(BUILD, name)
    src = os.path.join(EMUL, name + ".cu")
    deps = [src] + [os.path.join(ROOT, "cantera_b200", "csrc", f) for f in os.listdir(os.path.join(ROOT, "cantera_b200", "csrc"))]
    if not os.path.exists(exe) or any(os.path.getmtime(d) > os.path.getmtime(exe) for d in deps):
        subprocess.check_call([NVCC, "-std=c++17", "-O2", "-gencode", "arch=compute_100a,code=sm_100a", "-x", "cu", "-o", exe, src,
                               os.path.join(ROOT, "cantera_b200", "csrc", "mech_pack.cpp"),
                               "-diag-suppress", "177", "-Wno-deprecated-gpu-targets"])
    return exe


@pytest.fixture(scope="module")
def emul_thread():
    return _build("emul_thread")


@pytest.fixture(scope="module")
def emul_batch():
    return _build("emul_batch")


@pytest.mark.parametrize("case", ["h2o2", "gri30", "nDodecane", "pdep-test", "sri-falloff", "tsang-falloff",
                                  "jacobian-tests", "chemically-activated-reaction", "explicit-third-bodies"])
@pytest.mark.parametrize("unroll", [4, 2])
def test_thread_rates_vs_oracle(case, unroll, emul_thread, oracle_lib, tmp_path):
    out = str(tmp_path / "w.bin")
    subprocess.check_call([emul_thread, MECHS[case], "200", "11", out, str(unroll)], stderr=subprocess.DEVNULL)
    raw = open(out, "rb").read()
    n, K = np.frombuffer(raw, dtype="<i8", count=2)
    a = np.frombuffer(raw, dtype="<f8", offset=16)
    T, P = a[:n], a[n:2 * n]
    Y = a[2 * n:2 * n + n * K].reshape(n, K)
    W = a[2 * n + n * K:].reshape(n, K)
    o = oracle_lib.OracleMech(MECHS[case]).rates(T, P, Y, want=("wdot", "cdot", "ddot"))
    scale = o["cdot"] + o["ddot"]
    ok = np.isfinite(scale) & (scale > 0)
    err = np.abs(W - o["wdot"])[ok] / scale[ok]
    assert err.max() <= 1e-12


def _run_batch(exe, case, kind, nmax, slots, tmp_path):
    g = golden(kind, case)
    N, K, R = [int(x) for x in g["dims"]]
    n = min(N, nmax)
    tEnd = float(g["tEnd"][0])
    dT = float(g["dTign"][0]) if "dTign" in g else 0.0
    fin, fout = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    with open(fin, "wb") as f:
        f.write(struct.pack("<qqdd", n, K, tEnd, dT))
        f.write(np.ascontiguousarray(g["T"][:n]).tobytes())
        f.write(np.ascontiguousarray(g["P"][:n]).tobytes())
        f.write(np.ascontiguousarray(g["Y"].reshape(N, K)[:n]).tobytes())
    subprocess.check_call([exe, MECHS[case], fin, fout, str(slots)], stderr=subprocess.DEVNULL)
    a = np.frombuffer(open(fout, "rb").read(), dtype="<f8", offset=16)
    y = a[:n * (K + 1)].reshape(n, K + 1)
    tign = a[n * (K + 1):n * (K + 2)]
    status = a[n * (K + 2):n * (K + 3)]
    steps = a[n * (K + 3):]
    from cantera_b200.ctbm import read_arrays
    W = read_arrays(MECHS[case])["mol_weights"]
    Yo = y[:, 1:] * W
    Yo /= Yo.sum(axis=1, keepdims=True)
    return g, n, tEnd, y[:, 0], Yo, tign, status, steps


@pytest.mark.parametrize("case,nmax", [("h2o2", 32), ("gri30", 6)])
def test_batch_controller_ignition_vs_reference(case, nmax, emul_batch, tmp_path):
    g, n, tEnd, T, Y, tign, status, steps = _run_batch(emul_batch, case, "ignite", nmax, 3, tmp_path)
    N, K, R = [int(x) for x in g["dims"]]
    assert np.all(status == 0)
    tref = g["tign"][:n]
    ign = np.isfinite(tref)
    assert np.array_equal(np.isfinite(tign), ign)
    assert np.max(np.abs(tign[ign] - tref[ign]) / tref[ign]) <= 1e-3
    ok = ~(ign & (np.abs(tref - tEnd) <= 0.05 * tEnd))
    Tref, Yref = g["Tend"][:n], g["Yend"].reshape(N, K)[:n]
    assert np.max(np.abs(T[ok] - Tref[ok]) / Tref[ok]) <= 1e-6
    assert (np.abs(Y[ok] - Yref[ok]) - (1e-6 * np.abs(Yref[ok]) + 1e-12)).max() <= 0.0
    ref_steps = g["stats"].reshape(N, 8)[:n, 0].sum()
    assert 0.5 < steps.sum() / ref_steps < 2.0


def test_batch_controller_substep_vs_reference(emul_batch, tmp_path):
    g, n, tEnd, T, Y, tign, status, steps = _run_batch(emul_batch, "gri30", "substep", 8, 5, tmp_path)
    N, K, R = [int(x) for x in g["dims"]]
    assert np.all(status == 0)
    Tref, Yref = g["Tend"][:n], g["Yend"].reshape(N, K)[:n]
    assert np.max(np.abs(T - Tref) / Tref) <= 1e-6
    assert (np.abs(Y - Yref) - (1e-6 * np.abs(Yref) + 1e-12)).max() <= 0.0
