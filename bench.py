#!/usr/bin/env python
"""bench.py — FDDP OCP solves/s (batched, T=100) on B200, BASELINE.json's metric.

Workload at every N: BASELINE config C2 per GPU — SolverFDDP on ActionModelUnicycle (nx=3, nu=2, T=100),
65,536 random initial states per GPU (weak scaling: independent OCPs sharded over ranks, no data-path
collective; one NCCL all-gather of {cost, iter, stop, converged} at the end of every step).
A "step" = one batched solve() of all problems of the rank, run to each problem's own return.

  value : solves/s with init_xs/init_us already resident in HBM (cgx_solver_solve_resident), CUDA events on
          the solver's own stream, max over ranks.
  e2e   : the same through the reference-facing call with HOST buffers (cgx_solver_solve + xs/us/scalars
          accessors): pinned host -> device copies and device -> host reads inside the timed region.
  roofline     : the solve kernel's algorithmic HBM bytes (from the per-problem sweep counters) / its duration.
  cpu_baseline : the CPU oracle (restated reference algorithm, OpenMP over problems) on a bounded sample.

`--impl reference` times the reference algorithm's CPU implementation (the oracle port; the reference itself
cannot be compiled on this image — no Eigen/Boost) with all host threads on the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

T = 100
B_PER_GPU = 65536
NX, NU = 3, 2
METRIC = "fddp_ocp_solves_per_sec_T100"
UNIT = "solves/s"
# algorithmic HBM bytes per node of the lean storage policy (DESIGN.md "bytes per node"): useful doubles only,
# not the sector padding of the record layout
NL = NX * (NX + 1) // 2
BYTES_BWD_NODE = 8 * (NX + NU + NX + NU * NX + NU + NX + NX + NL)  # read xs,us ; write fs,K,k,w,Vx,tril(Vxx)
BYTES_FWD_NODE = 8 * (NX + NU + NU * NX + NU + NX + NX + NX + NU)  # read xs,us,K,k,fs,w ; write xs_try,us_try
BYTES_CALC_NODE = 8 * (NX + NU)


def workload(rank, B):
    rng = np.random.default_rng(1000 * rank)  # seed 0 on rank 0
    x0 = np.stack([rng.uniform(-2, 2, B), rng.uniform(-2, 2, B), rng.uniform(-np.pi, np.pi, B)], 1)
    xs = np.repeat(x0[:, None, :], T + 1, 1)
    us = np.zeros((B, T, NU))
    return x0, xs, us


class ClockSampler(object):
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for i, n in enumerate(names):
                if len(r) > 3 + i and r[3 + i].lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_oracle_rate(nsample, nthreads=0, reps=1):
    """Reference algorithm on the host cores: oracle, -O3 build, OpenMP over problems."""
    import oracle_lib as ol

    x0, xs, us = workload(0, B_PER_GPU)
    x0, xs, us = x0[:nsample], xs[:nsample], us[:nsample]
    ob = ol.OracleBatch("unicycle", "fddp", T, NX, NU, ol.unicycle_params(), x0, fast=True)
    best = None
    for _ in range(reps):
        ob.solve(xs, us, 100, False, 1e-9, nthreads=nthreads)
        best = ob.seconds if best is None else min(best, ob.seconds)
    cores = ob.lib.orc_max_threads() if nthreads <= 0 else nthreads
    return nsample / best, cores, best


def run_reference(args, rank, world):
    if rank != 0:
        return
    import __graft_entry__ as ge

    ge.build(quiet=True)
    nsample = 8192
    times = []
    import oracle_lib as ol

    x0, xs, us = workload(0, B_PER_GPU)
    x0, xs, us = x0[:nsample], xs[:nsample], us[:nsample]
    ob = ol.OracleBatch("unicycle", "fddp", T, NX, NU, ol.unicycle_params(), x0, fast=True)
    cores = ob.lib.orc_max_threads()
    for i in range(args.warmup + args.steps):
        ob.solve(xs, us, 100, False, 1e-9, nthreads=0)
        if i >= args.warmup:
            times.append(ob.seconds)
    tot = sum(times)
    value = nsample * len(times) / tot
    sample = "first %d of the %d problems of config C2 per step (same seeds), OpenMP over problems" % (nsample, B_PER_GPU)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * tot / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C2: SolverFDDP, ActionModelUnicycle nx=3 nu=2 T=100, random x0, maxiter=100",
                   "problems_per_step": nsample, "note": "CPU port of the reference algorithm (oracle/); the reference "
                   "itself needs Eigen+Boost which this image lacks"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--batch", type=int, default=B_PER_GPU)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import __graft_entry__ as ge

    if rank == 0:
        ge.build(quiet=True)
    dist = None
    if world > 1:
        import torch.distributed as dist

        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist.barrier()
    import crocoddyl_b200 as cb

    cb.set_device(local_rank)
    torch.cuda.set_device(local_rank)
    B = args.batch
    x0, xs_np, us_np = workload(rank, B)
    # pinned host buffers (what a caller hands to the reference-facing API)
    xs_h = torch.from_numpy(xs_np).pin_memory()
    us_h = torch.from_numpy(us_np).pin_memory()
    xs_out = torch.empty_like(xs_h).pin_memory()
    us_out = torch.empty_like(us_h).pin_memory()
    model = cb.ActionModelUnicycle()
    problem = cb.ShootingProblem(x0, [model] * T, model)
    solver = cb.SolverFDDP(problem)
    solver.setStorage(False)  # lean: keeps xs, us, K, k, Vx, Vxx, fs (the accessors north_star names)
    L = cb.lib()
    h = solver._h
    conv = np.zeros(B, dtype=np.uint8)
    scal = np.zeros((B, 17))
    import ctypes as C

    def dp(t):
        return C.cast(t.data_ptr(), C.POINTER(C.c_double))

    def gather_results():
        # the one collective of the path: per-problem {cost, iter, stop, converged} to every rank
        if dist is None:
            return None
        rec = torch.from_numpy(np.ascontiguousarray(scal[:, [0, 9, 1, 11]])).cuda()
        out = torch.empty((world,) + tuple(rec.shape), dtype=rec.dtype, device=rec.device)
        dist.all_gather_into_tensor(out, rec)
        return out

    def step_resident():
        cb.check(L.cgx_solver_solve_resident(h, 100, 0, 1e-9))
        cb.check(L.cgx_solver_download(h, conv.ctypes.data_as(C.POINTER(C.c_ubyte))))
        cb.check(L.cgx_solver_get_scalars(h, cb.dptr(scal)))
        gather_results()
        return solver.timing()

    def step_e2e():
        cb.check(L.cgx_solver_solve(h, dp(xs_h), dp(us_h), 100, 0, 1e-9, conv.ctypes.data_as(C.POINTER(C.c_ubyte))))
        cb.check(L.cgx_solver_get_all(h, b"xs", dp(xs_out)))
        cb.check(L.cgx_solver_get_all(h, b"us", dp(us_out)))
        cb.check(L.cgx_solver_get_scalars(h, cb.dptr(scal)))
        gather_results()

    def sync():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- resident (value) -------------------------------------------------------------------
    cb.check(L.cgx_solver_upload(h, dp(xs_h), dp(us_h)))
    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    sync()
    call_ms, kern_ms, launches = [], [], 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        tm = step_resident()
        call_ms.append(tm["call_ms"])
        kern_ms.append(tm["kernel_ms"])
        launches += tm["launches"] + 1  # + the scalar export kernel
    sync()
    wall_ms = 1e3 * (time.perf_counter() - t0)
    clocks = sampler.stop() if rank == 0 else None
    dev_ms = float(np.sum(call_ms))
    sg = solver.scalars()
    nb, nf = sg["n_backward"], sg["n_forward"]
    alg_bytes = float(T * (nb * BYTES_BWD_NODE + (nf - 1) * BYTES_FWD_NODE + BYTES_CALC_NODE).sum())
    # ---- e2e ---------------------------------------------------------------------------------
    for _ in range(2):
        step_e2e()
    sync()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    sync()
    e2e_ms = 1e3 * (time.perf_counter() - t0)
    if dist is not None:
        v = torch.tensor([dev_ms, e2e_ms, wall_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        dev_ms, e2e_ms, wall_ms = [float(a) for a in v.cpu()]
    if rank == 0:
        total = B * world * args.steps
        value = total / (dev_ms * 1e-3)
        peak, peak_src = hbm_peak()
        kms = float(np.mean(kern_ms))
        achieved = alg_bytes / (kms * 1e-3) / 1e9
        h2d = int(xs_h.numel() * 8 + us_h.numel() * 8)
        d2h = int(xs_out.numel() * 8 + us_out.numel() * 8 + scal.size * 8 * 2 + B)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C2: SolverFDDP, ActionModelUnicycle nx=3 nu=2 T=100, %d random x0 per GPU, "
                       "init_xs={x0}x101, init_us=0, maxiter=100, reginit=1e-9" % B,
                       "problems_per_gpu": B, "storage": "lean (xs,us,K,k,Vx,Vxx,fs kept)",
                       "l2": "working set 1.9 GB per GPU >> 126 MB L2 (inputs larger than L2)",
                       "parallelism": "independent OCPs sharded over %d rank(s); NCCL all-gather of per-problem "
                       "{cost,iter,stop,converged}" % world},
            "e2e": {"value": total / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / args.steps},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "wall_ms_per_step": wall_ms / args.steps,
            "solve_stats": {"iterations_mean": float(sg["n_records"].mean()), "converged_frac": float(conv.mean()),
                            "backward_sweeps_mean": float(nb.mean()), "forward_sweeps_mean": float(nf.mean())},
            "roofline": {"kernel": "tpp_solve_kernel<UnicycleModel>", "bound": "hbm", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": kms},
        }
        if not args.no_cpu_baseline:
            nsample = 8192
            rate, cores, sec = cpu_oracle_rate(nsample)
            line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "first %d problems of the same workload, %.2f s, OpenMP over problems" %
                                    (nsample, sec)}
        prof = os.path.join(ROOT, "profiles", "traffic_r01.json")
        if os.path.exists(prof):
            try:
                line["roofline"]["traffic"] = json.load(open(prof)).get("tpp_solve_kernel_dram_bytes_per_launch")
            except Exception:
                pass
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
