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


if __name__ == "__main__":
    # bench_riccati.py [n ...] [--B batch] [--T horizon]: development knobs for profiling one size on a small batch
    argv = sys.argv[1:]
    kw = {}
    if "--B" in argv:
        kw["B"] = int(argv[argv.index("--B") + 1])
    if "--T" in argv:
        kw["T"] = int(argv[argv.index("--T") + 1])
    sizes = tuple(int(a) for i, a in enumerate(argv) if a.isdigit() and (i == 0 or argv[i - 1] not in ("--B", "--T")))
    print(json.dumps(run(sizes or (12, 36, 72), **kw)))
