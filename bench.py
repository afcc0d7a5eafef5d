#!/usr/bin/env python
"""bench.py — the headline measurement (BASELINE.json): GRI-30 net-production-rate evaluations/s
per B200 (config[1], 2^22 random states), plus 1 ms constant-pressure ignition reactors/s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload wdot|ignition]

One "step" = one pass of the hot path over the whole batch. States have no inter-GPU dependence:
under torchrun every rank owns its own 2^22-state batch (weak scaling, no collective on the data
path); timing is CUDA events on the launching stream, max over ranks.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GRI = dict(K=53, R=325)
FLOP_PER_STATE = 38374.0        # algorithmic FP64 flop per gri30 wdot evaluation
BYTES_PER_STATE = 8 * (53 + 2) + 8 * 53  # 864 B: T, P, Y in; wdot out
FP64_PEAK_TFLOPS_NOMINAL = 148 * 64 * 2 * 1.965e9 / 1e12  # 37.2: no FP64 entry in MEASURED_PEAKS.json
WDOT_TRAFFIC_BYTES_PER_STATE = 874.8  # dram read+write per state, ncu --set full (profiles/r01_rates_kernel*.txt)
# ignition: algorithmic FP64 flop per reactor from the solver's own counters:
# n_rhs (F_wdot + 6K) + n_setup (2/3 (K+1)^3 + F_jac) + n_newton 2 (K+1)^2, F_jac ~ 1.5 F_wdot
def ignition_flops(stats, K=53):
    n = K + 1
    st = stats.reshape(-1, 8).sum(axis=0).astype(float)
    rhs, setups, nni = st[1], st[3], st[4]
    return rhs * (FLOP_PER_STATE + 6 * K) + setups * (2.0 / 3.0 * n ** 3 + 1.5 * FLOP_PER_STATE) + nni * 2.0 * n * n


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="wdot", choices=["wdot", "ignition"])
    ap.add_argument("--states", type=int, default=1 << 22)
    ap.add_argument("--reactors", type=int, default=1 << 17)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ignition", action="store_true", help="wdot workload: skip the short ignition run reported under 'ignition'")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons with nvidia-smi while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.stop_flag = False
        self.sm, self.reasons, self.sm_max = [], set(), None

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_throttle_reasons.hw_slowdown,clocks_throttle_reasons.hw_thermal_slowdown," \
            "clocks_throttle_reasons.sw_thermal_slowdown,clocks_throttle_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                self.sm.append(float(f[0]))
                self.sm_max = float(f[1])
                for nm, v in zip(names, f[2:]):
                    if v.lower().startswith("active"):
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.05)

    def result(self):
        s = sorted(self.sm)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons)}


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ------------------------------------------------------------------------------ reference arm

REF_BENCH = os.path.join(ROOT, "oracle", "_ref", "ref_bench")
REF_DATA = os.path.join(ROOT, "oracle", "_ref", "data")
IGN_ARGS = ["CH4", "2", "O2", "N2", "1000", "2000", "1", "50", "1e-3"]


def run_reference(workload, n, procs, seed=20261104):
    """The reference's own CPU implementation on `n` states over `procs` worker processes.
    oracle/_ref/ref_bench (the unmodified reference compiled from its sources) when present
    ("reference"), else the plain-C oracle port on one core ("port")."""
    if os.path.exists(REF_BENCH):
        mode = "wdot" if workload == "wdot" else "ignite"
        cmd = [REF_BENCH, mode, "gri30.yaml", "gri30", str(n), str(seed), str(procs), REF_DATA]
        if mode == "ignite":
            cmd += IGN_ARGS
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=1800)
        line = [l for l in out.stdout.splitlines() if l.startswith("{")]
        if out.returncode == 0 and line:
            r = json.loads(line[-1])
            return {"kind": "reference", "cores": procs, "seconds": r["seconds"], "per_second": r["per_second"], "n": n}
    if workload != "wdot":
        return None
    import numpy as np
    from oracle.oracle_py import OracleMech
    m = OracleMech(os.path.join(ROOT, "cantera_b200", "data", "gri30.ctbm"))
    rng = np.random.default_rng(seed)
    T = rng.uniform(800, 2500, n)
    P = 101325.0 * rng.uniform(1, 50, n)
    Y = rng.uniform(0, 1, (n, m.K))
    Y /= Y.sum(axis=1, keepdims=True)
    t0 = time.perf_counter()
    m.rates(T, P, Y, want=("wdot",))
    dt = time.perf_counter() - t0
    return {"kind": "port", "cores": 1, "seconds": dt, "per_second": n / dt, "n": n}


def metric_of(workload):
    if workload == "wdot":
        return "gri30_wdot_evals_per_s", "states/s", "gri30 getNetProductionRates, random (T 800-2500 K, P 1-50 atm, random Y)"
    return "gri30_1ms_ignition_reactors_per_s", "reactors/s", \
        "gri30 IdealGasConstPressureMoleReactor advance(1e-3), CH4/air phi 0.5-2, T0 1000-2000 K, P 1-50 atm"


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    metric, unit, desc = metric_of(args.workload)
    cores = host_cores()
    # bounded sample: ~3 s (wdot) / ~15 s (ignition) of CPU work per step
    n = 120000 * cores if args.workload == "wdot" else 24 * cores
    for _ in range(args.warmup):
        run_reference(args.workload, max(cores, n // 8), cores)
    secs, total = 0.0, 0
    kind = "reference"
    for _ in range(args.steps):
        r = run_reference(args.workload, n, cores)
        if r is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_bench missing for the ignition workload"}))
            return 0
        secs += r["seconds"]
        total += r["n"]
        kind, cores_used = r["kind"], r["cores"]
    value = total / secs
    sample = "%d states per step on %d host processes (one Solution%s per process)" % (
        n, cores_used, "" if args.workload == "wdot" else " + ReactorNet")
    line = {
        "impl": "reference", "metric": metric, "value": value, "unit": unit, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "states_per_step": n},
        "cpu_baseline": {"value": value, "unit": unit, "cores": cores_used, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------ our arm

def make_states(torch, n, K, seed, device):
    gen = torch.Generator(device=device).manual_seed(seed)
    T = 800 + 1700 * torch.rand(n, dtype=torch.float64, device=device, generator=gen)
    P = 101325.0 * (1 + 49 * torch.rand(n, dtype=torch.float64, device=device, generator=gen))
    Y = torch.rand(n, K, dtype=torch.float64, device=device, generator=gen)
    Y /= Y.sum(dim=1, keepdim=True)
    return T, P, Y


def make_ignition_states(torch, n, seed, device):
    """CH4/air, phi ~ U(0.5, 2), T0 ~ U(1000, 2000) K, P ~ U(1, 50) atm."""
    import numpy as np
    from cantera_b200.ctbm import read_arrays
    tab = read_arrays(os.path.join(ROOT, "cantera_b200", "data", "gri30.ctbm"))
    names = tab["species_names"].split("\n")
    W = tab["mol_weights"]
    rng = np.random.default_rng(seed)
    T = rng.uniform(1000, 2000, n)
    P = 101325.0 * rng.uniform(1, 50, n)
    phi = rng.uniform(0.5, 2.0, n)
    X = np.zeros((n, len(names)))
    X[:, names.index("CH4")] = phi
    X[:, names.index("O2")] = 2.0
    X[:, names.index("N2")] = 7.52
    Y = X * W
    Y /= Y.sum(axis=1, keepdims=True)
    return T, P, Y


def ours(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this engine has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from cantera_b200 import BatchedGasKinetics, BatchReactorNet
    kin = BatchedGasKinetics("gri30", device=local)
    kin.use_torch_stream()
    K = kin.n_species
    metric, unit, desc = metric_of(args.workload)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    launches = 0
    # FP64 roofline denominator: measured live by the library's DFMA micro-benchmark (MEASURED_PEAKS.json
    # carries no FP64 figure); the nominal number is kept beside it
    from cantera_b200 import clib
    fp64_peak = float(clib.lib().ctb_fp64PeakTflops(local))
    if not fp64_peak > 0:
        fp64_peak = FP64_PEAK_TFLOPS_NOMINAL
    ign_stats = {}
    if args.workload == "wdot":
        n = args.states
        T, P, Y = make_states(torch, n, K, 20261104 + rank, dev)
        out = torch.empty(n, K, dtype=torch.float64, device=dev)
        hT, hP, hY = T.cpu().pin_memory(), P.cpu().pin_memory(), Y.cpu().pin_memory()
        hout = torch.empty(n, K, dtype=torch.float64).pin_memory()

        def step():
            kin.net_production_rates(T, P, Y, out=out)

        def step_e2e():
            kin.net_production_rates(hT.numpy(), hP.numpy(), hY.numpy(), out=hout.numpy())

        units, h2d, d2h, launches_per_step = n, 8 * n * (K + 2), 8 * n * K, 1
        l2_note = "inputs+outputs 3.6 GB per step >> 126 MB L2"
    else:
        n = args.reactors
        Tn, Pn, Yn = make_ignition_states(torch, n, 20261104 + rank, dev)
        T, P, Y = (torch.from_numpy(a).to(dev) for a in (Tn, Pn, Yn))
        net = BatchReactorNet(kin, n)
        state = {}

        def step():
            net.set_state(T, P, Y)
            net.advance(1e-3)
            state["out"] = net.get_state(device_like=T)
            ign_stats["stats"] = net.solver_stats()

        def step_e2e():
            net.set_state(Tn, Pn, Yn)
            net.advance(1e-3)
            state["host"] = net.get_state()

        units, h2d, d2h, launches_per_step = n, 8 * n * (K + 2), 8 * n * (K + 1), 0
        l2_note = "reactor state and integrator memory %d MB per step" % (n * 8 * 12 * (K + 1) >> 20)

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    per_step = []
    ev0.record()
    for _ in range(args.steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step()
        b.record()
        per_step.append((a, b))
        launches += launches_per_step
        if args.workload == "ignition":
            launches += 3 + 4 * net.last_rounds()  # init, contents + (memset-free) control, rhs, setup, lu per round
    ev1.record()
    barrier()
    sampler.stop_flag = True
    sampler.join()
    ms = ev0.elapsed_time(ev1)
    kernel_ms = sum(a.elapsed_time(b) for a, b in per_step) / args.steps
    # end to end through the C ABI with HOST buffers (pinned), copies inside the timed region
    e2e_steps = max(1, min(args.steps, 3))
    step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    torch.cuda.synchronize()
    e2e_ms = 1e3 * (time.perf_counter() - t0)
    if world > 1:
        t = torch.tensor([ms, e2e_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_ms = t.tolist()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0
    value = units * world * args.steps / (ms * 1e-3)
    e2e_value = units * world * e2e_steps / (e2e_ms * 1e-3)
    peaks = measured_peaks()
    line = {
        "metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "states_per_gpu": units, "mechanism": "gri30 (53 species, 325 reactions)",
                   "l2": l2_note},
        "clocks": sampler.result(),
        "e2e": {"value": e2e_value, "unit": unit, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps},
        "gpu_launches": launches,
    }
    if args.workload == "wdot":
        # dominant kernel = ratesKernel (the only launch of a step). FP64-pipe bound:
        # the contract's "tensor" slot carries the compute roofline; the HBM view is given alongside.
        tflops = FLOP_PER_STATE * units / (kernel_ms * 1e-3) / 1e12
        gbs = BYTES_PER_STATE * units / (kernel_ms * 1e-3) / 1e9
        line["roofline"] = {
            "bound": "tensor", "achieved": tflops, "peak": fp64_peak, "unit": "TFLOP/s",
            "frac": tflops / fp64_peak, "traffic": WDOT_TRAFFIC_BYTES_PER_STATE * units,
            "note": "FP64 DFMA pipe (no tensor cores on this path; the contract's 'tensor' slot carries the compute "
                    "roofline). peak = measured live by ctb_fp64PeakTflops (register-resident DFMA micro-benchmark; "
                    "nominal 148 SM x 64 DFMA/clk x 1.965 GHz = %.1f); algorithmic flop/state = 38374 (SURVEY.md 8(d)); "
                    "traffic = dram bytes/state of the ncu capture in profiles/ x states" % FP64_PEAK_TFLOPS_NOMINAL,
            "hbm": {"achieved_gbs": gbs, "peak_gbs": peaks.get("hbm_gbs"), "frac": gbs / peaks["hbm_gbs"] if peaks.get("hbm_gbs") else None,
                    "bytes_per_state": BYTES_PER_STATE},
            "kernel_ms": kernel_ms,
        }
    if args.workload == "ignition" and "stats" in ign_stats:
        import numpy as np
        st = np.asarray(ign_stats["stats"]).reshape(-1, 8)
        fl = ignition_flops(st)
        tflops = fl / (kernel_ms * 1e-3) / 1e12
        tot = st.sum(axis=0)
        line["roofline"] = {
            "bound": "tensor", "achieved": tflops, "peak": fp64_peak, "unit": "TFLOP/s", "frac": tflops / fp64_peak,
            "traffic": None,
            "note": "FP64 DFMA pipe; algorithmic flop from the solver's counters: n_rhs (F_wdot + 6K) + n_setup "
                    "(2/3 (K+1)^3 + 1.5 F_wdot) + n_newton 2 (K+1)^2; peak measured live by ctb_fp64PeakTflops",
            "per_reactor": {"steps": tot[0] / n, "rhs": tot[1] / n, "setups": tot[3] / n, "newton_iters": tot[4] / n},
            "rounds": net.last_rounds(), "kernel_ms": kernel_ms,
        }
    if args.workload == "wdot" and not args.no_ignition and world == 1:
        # the second half of BASELINE.json's metric, on a bounded batch (one timed advance after one warm-up)
        n_ign = 131072
        Tn, Pn, Yn = make_ignition_states(torch, n_ign, 20261104, dev)
        Ti, Pi, Yi = (torch.from_numpy(a).to(dev) for a in (Tn, Pn, Yn))
        net = BatchReactorNet(kin, n_ign)
        for rep_i in range(2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            net.set_state(Ti, Pi, Yi)
            net.advance(1e-3)
            torch.cuda.synchronize()
            ign_s = time.perf_counter() - t0
        line["ignition"] = {"metric": "gri30_1ms_ignition_reactors_per_s", "value": n_ign / ign_s, "unit": "reactors/s",
                            "reactors": n_ign, "seconds": ign_s, "failed": int((net.status() != 0).sum())}
        net.close()
    if not args.no_cpu_baseline:
        cores = host_cores()
        nref = 400000 * cores if args.workload == "wdot" else 24 * cores
        r = run_reference(args.workload, nref, cores)
        if r:
            line["cpu_baseline"] = {"value": r["per_second"], "unit": unit, "cores": r["cores"], "kind": r["kind"],
                                    "sample": "%d states, %.1f s wall" % (r["n"], r["seconds"])}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    a = parse()
    sys.exit(reference_arm(a) if a.impl == "reference" else ours(a))
