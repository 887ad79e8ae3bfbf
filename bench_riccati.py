#!/usr/bin/env python
"""Riccati roofline study (BASELINE config C3): one SolverDDP::backwardPass (computeDirection(false)) with the
derivatives resident, batched LQR nx in {12,36,72}, nu = nx/3, T = 100, B = 4096. Reports us per problem-sweep
and FP64 FLOP/s against the measured DFMA peak (cgx_probe_fp64_tflops). Flop count = literal operation
sequence of ddp.cpp:202-244: F_ric(n,m) = 4n^3 + 6n^2 m + 4n m^2 + m^3/3 + 6n^2 + 6nm + 4m^2 per node."""
import json
import sys
import os

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def f_ric(n, m):
    return 4 * n ** 3 + 6 * n * n * m + 4 * n * m * m + m ** 3 / 3.0 + 6 * n * n + 6 * n * m + 4 * m * m


def random_lqr(cb, n, m, B, seed):
    rng = np.random.default_rng(seed)
    mod = cb.ActionModelLQR(n, m, True)
    mod.Fx = np.eye(n) + 0.1 * rng.standard_normal((B, n, n)) / np.sqrt(n)
    mod.Fu = rng.standard_normal((B, n, m))
    A = rng.standard_normal((B, n, n))
    Bm = rng.standard_normal((B, m, m))
    mod.Lxx = np.einsum("bki,bkj->bij", A, A) / n + np.eye(n)
    mod.Luu = np.einsum("bki,bkj->bij", Bm, Bm) / m + np.eye(m)
    mod.Lxu = 0.1 * rng.standard_normal((B, n, m))
    mod.lx = rng.standard_normal((B, n))
    mod.lu = rng.standard_normal((B, m))
    return mod, rng.standard_normal((B, n))


def run(sizes=(12, 36, 72), B=4096, T=100, reps=3, peak_tflops=None):
    import crocoddyl_b200 as cb

    if peak_tflops is None:
        peak_tflops = cb.probe_fp64_tflops()
    out = []
    for n in sizes:
        m = n // 3
        mod, x0 = random_lqr(cb, n, m, B, 1)
        problem = cb.ShootingProblem(x0, [mod] * T, mod)
        solver = cb.SolverFDDP(problem)
        solver.setStorage(False)
        xs = np.repeat(x0[:, None, :], T + 1, 1)
        solver.setCandidate(xs, np.zeros((B, T, m)), False)
        solver.x_reg = solver.u_reg = 1e-9
        solver.computeDirection(True)  # calc + calcDiff + gaps resident, warm-up
        ms = []
        for _ in range(reps):
            solver.computeDirection(False)  # = backwardPass()
            ms.append(solver.timing()["kernel_ms"])
        best = min(ms)
        flops = f_ric(n, m) * T * B
        tf = flops / (best * 1e-3) / 1e12
        out.append({"nx": n, "nu": m, "T": T, "B": B, "sweep_ms_batched": best, "us_per_problem_sweep": 1e3 * best / B,
                    "fp64_tflops": tf, "fp64_peak_tflops_measured_dfma": peak_tflops, "frac": tf / peak_tflops})
        del solver, problem
    return out


def run_c1(trials=2000, cpu_trials=1000):
    """BASELINE config C1 = benchmark/lqr-optctrl.cpp:16-54: ONE OCP, ActionModelLQR(37, 12) defaults, T = 100, x0 = 0,
    xs = {x0} x 101, us = {0} x 100, SolverDDP::solve(xs, us, MAXITER = 1), timed per call like the reference prints it
    (mean (min-max) in ms over the trials; the reference uses 5000). GPU: through the reference-facing solve with host
    buffers (H2D + one-CTA kernel + D2H; a single OCP is a latency measurement, the batch is the product). CPU: the
    oracle port on one thread (the reference binary cannot be built on this box)."""
    import time
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests"))
    import crocoddyl_b200 as cb
    import oracle_lib as ol

    n, m, T = 37, 12, 100
    x0 = np.zeros((1, n))
    xs, us = np.zeros((1, T + 1, n)), np.zeros((1, T, m))
    mod = cb.ActionModelLQR(n, m, True)
    problem = cb.ShootingProblem(x0, [mod] * T, mod)
    solver = cb.SolverDDP(problem)

    def stats(v):
        v = np.asarray(v)
        return {"mean_ms": float(v.mean()), "min_ms": float(v.min()), "max_ms": float(v.max()), "trials": int(v.size)}

    for _ in range(20):
        solver.solve(xs, us, 1)
    call, kern = [], []
    for _ in range(trials):
        t0 = time.perf_counter()
        solver.solve(xs, us, 1)
        call.append(1e3 * (time.perf_counter() - t0))
        kern.append(solver.timing()["kernel_ms"])
    ob = ol.OracleBatch("lqr", "ddp", T, n, m, ol.lqr_params(n, m, True), x0, fast=True)
    cpu = []
    for _ in range(20):
        ob.solve(xs, us, 1, False, 1e-9, 1)
    for _ in range(cpu_trials):
        ob.solve(xs, us, 1, False, 1e-9, 1)
        cpu.append(1e3 * ob.seconds)
    np.testing.assert_allclose(solver.get("xs")[0], ob.get_all("xs")[0], rtol=1e-9, atol=1e-9)
    return {"config": "C1: SolverDDP::solve(xs, us, 1) on ActionModelLQR(37,12), T=100, single OCP "
                      "(benchmark/lqr-optctrl.cpp)", "gpu_call": stats(call), "gpu_kernel": stats(kern),
            "cpu_oracle_1thread": dict(stats(cpu), kind="port")}


if __name__ == "__main__":
    if "--c1" in sys.argv:
        print(json.dumps(run_c1()))
        sys.exit(0)
    # bench_riccati.py [n ...] [--B batch] [--T horizon]: development knobs for profiling one size on a small batch
    argv = sys.argv[1:]
    kw = {}
    if "--B" in argv:
        kw["B"] = int(argv[argv.index("--B") + 1])
    if "--T" in argv:
        kw["T"] = int(argv[argv.index("--T") + 1])
    sizes = tuple(int(a) for i, a in enumerate(argv) if a.isdigit() and (i == 0 or argv[i - 1] not in ("--B", "--T")))
    print(json.dumps(run(sizes or (12, 36, 72), **kw)))
