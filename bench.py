#!/usr/bin/env python
"""bench.py — FDDP OCP solves/s (batched, T=100) on B200, BASELINE.json's metric.

Default workload at every N: BASELINE config C2 per GPU — SolverFDDP on ActionModelUnicycle (nx=3, nu=2, T=100),
65,536 random initial states per GPU (weak scaling: independent OCPs sharded over ranks, no data-path collective;
one NCCL all-gather of {cost, iter, stop, converged} at the end of every step).
`--workload arm` runs configs C4/C5 instead: the synthetic 7-DoF arm (Euler o FreeFwdDynamics), B=1024 per GPU, or
`--batch-total 8192` split contiguously over the ranks (strong scaling, config C5).
A "step" = one batched solve() of all problems of the rank, run to each problem's own return.

  value : solves/s with init_xs/init_us already resident in HBM (cgx_solver_solve_resident), CUDA events on
          the solver's own stream, max over ranks.
  e2e   : the same through the reference-facing call with HOST buffers (cgx_solver_solve + xs/us/scalars
          accessors): pinned host -> device copies and device -> host reads inside the timed region.
  roofline     : the solve kernel's algorithmic HBM bytes (from the per-problem sweep counters) / its duration.
  cpu_baseline : the CPU oracle (restated reference algorithm, OpenMP over problems) on a bounded sample.

`--impl reference` times the reference algorithm's CPU implementation (the oracle port; the reference itself
cannot be compiled on this image — no Eigen/Boost/Pinocchio) with all host threads on the same workload.
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
METRIC = "fddp_ocp_solves_per_sec_T100"
UNIT = "solves/s"


class Unicycle(object):
    """BASELINE config C2: ActionModelUnicycle defaults, x0 ~ U([-2,2]^2 x [-pi,pi]), seed 0."""
    name, nx, nu, reginit = "unicycle", 3, 2, 1e-9
    default_batch, cpu_sample = 65536, 8192
    kernel = "tpp_solve_kernel<UnicycleModel>"
    oracle_kind = "unicycle"
    # algorithmic HBM bytes per node of the lean storage policy: useful doubles only
    bytes_bwd = 8 * (3 + 2 + 3 + 2 * 3 + 2 + 3 + 3 + 6)  # read xs,us ; write fs,K,k,w,Vx,tril(Vxx)
    bytes_fwd = 8 * (3 + 2 + 2 * 3 + 2 + 3 + 3 + 3 + 2)  # read xs,us,K,k,fs,w ; write xs_try,us_try
    bytes_calc = 8 * (3 + 2)

    def describe(self, B):
        return ("C2: SolverFDDP, ActionModelUnicycle nx=3 nu=2 T=100, %d random x0 per GPU, init_xs={x0}x101, "
                "init_us=0, maxiter=100, reginit=1e-9" % B)

    def model(self):
        import crocoddyl_b200 as cb

        return cb.ActionModelUnicycle()

    def terminal(self, model):
        return model

    def oracle_params(self):
        import oracle_lib as ol

        return ol.unicycle_params()

    def inputs(self, lo, hi, total, seed):
        rng = np.random.default_rng(seed)
        x0 = np.stack([rng.uniform(-2, 2, total), rng.uniform(-2, 2, total), rng.uniform(-np.pi, np.pi, total)], 1)[lo:hi]
        return x0, np.repeat(x0[:, None, :], T + 1, 1), np.zeros((hi - lo, T, 2))


class Arm(object):
    """BASELINE configs C4/C5: synthetic 7-DoF arm, Euler(1e-3) o FreeFwdDynamics, costs gripperPose 1,
    xReg 1e-4, uReg 1e-4; q ~ U(-1,1)^7, v = 0, seed 2; init_us = quasiStatic(x0); reginit 0.1."""
    name, nx, nu, reginit = "arm", 14, 7, 0.1
    default_batch, cpu_sample = 1024, 256
    kernel = ("cta_phase_{nodes,direction,forward}_kernel<ArmModel<7>>: one solve = lockstep rounds of calc/calcDiff, "
              "backward pass and line-search launches over the whole batch (kernel_ms = the whole sequence)")
    oracle_kind = "arm"
    n, m = 14, 7
    # FP64 work per node: Riccati F_ric(14,7); calc / calcDiff = static operation count of the node functions
    # (tools/count_fp64_ops.py: DFMA = 2, DADD = DMUL = 1 flop)
    flops_ric = 4 * n**3 + 6 * n * n * m + 4 * n * m * m + m**3 / 3. + 6 * n * n + 6 * n * m + 4 * m * m
    flops_calc, flops_calcdiff = 7375., 22173.
    # CTA-per-problem core: node derivatives are written by calcDiff and read back by the backward pass
    bytes_bwd = 8 * ((n + m) + 2 * (2 * n * n + n * m + m * m + n + m) + n + m * n + m + 2 * n + n * n)
    bytes_fwd = 8 * (n + m + m * n + m + n + n + n + m + n + 1)
    bytes_calc = 8 * (n + m + n + 1)

    def describe(self, B):
        return ("C4/C5: SolverFDDP, 7-DoF synthetic arm (IntegratedActionModelEuler dt=1e-3 o DifferentialActionModel"
                "FreeFwdDynamics, gripperPose+xReg+uReg), nx=14 nu=7 T=100, %d random q0 per GPU, init_xs={x0}x101, "
                "init_us=quasiStatic(x0), maxiter=100, reginit=0.1" % B)

    def model(self):
        import crocoddyl_b200 as cb

        return cb.ArmActionModel(cb.synthetic_arm(7), 1e-3, w_gripperPose=1., w_xReg=1e-4, w_uReg=1e-4)

    def terminal(self, model):
        return model.terminal_model

    def oracle_params(self):
        return self.model().params()

    def inputs(self, lo, hi, total, seed):
        rng = np.random.default_rng(2 + seed)
        x0 = np.concatenate([rng.uniform(-1, 1, (total, 7)), np.zeros((total, 7))], 1)[lo:hi]
        return x0, np.repeat(x0[:, None, :], T + 1, 1), None  # us = quasiStatic(x0), filled by the caller


WORKLOADS = {"unicycle": Unicycle, "arm": Arm}


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


def oracle_inputs(wl, nsample):
    """First nsample problems of rank 0's workload, for the CPU legs."""
    import oracle_lib as ol

    x0, xs, us = wl.inputs(0, nsample, max(nsample, wl.default_batch), 0)
    params = wl.oracle_params()
    if us is None:
        u0 = np.stack([ol.arm_op(params, "quasistatic", x0[b, :wl.nu], wl.nu) for b in range(nsample)])
        us = np.repeat(u0[:, None, :], T, 1)
    return params, x0, xs, us


def host_threads():
    """All host threads this process may use (torchrun exports OMP_NUM_THREADS=1: ask the affinity mask instead)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_oracle_rate(wl, nsample, nthreads=0, reps=1):
    """Reference algorithm on the host cores: oracle, -O3 build, OpenMP over problems."""
    import oracle_lib as ol

    nthreads = nthreads or host_threads()

    params, x0, xs, us = oracle_inputs(wl, nsample)
    ob = ol.OracleBatch(wl.oracle_kind, "fddp", T, wl.nx, wl.nu, params, x0, fast=True)
    best = None
    for _ in range(reps):
        ob.solve(xs, us, 100, False, wl.reginit, nthreads=nthreads)
        best = ob.seconds if best is None else min(best, ob.seconds)
    cores = ob.lib.orc_max_threads() if nthreads <= 0 else nthreads
    return nsample / best, cores, best


def run_reference(args, wl, rank, world):
    if rank != 0:
        return
    import __graft_entry__ as ge

    ge.build(quiet=True)
    import oracle_lib as ol

    nsample = wl.cpu_sample
    params, x0, xs, us = oracle_inputs(wl, nsample)
    ob = ol.OracleBatch(wl.oracle_kind, "fddp", T, wl.nx, wl.nu, params, x0, fast=True)
    cores = host_threads()
    times = []
    for i in range(args.warmup + args.steps):
        ob.solve(xs, us, 100, False, wl.reginit, nthreads=cores)
        if i >= args.warmup:
            times.append(ob.seconds)
    tot = sum(times)
    value = nsample * len(times) / tot
    sample = "first %d problems of the workload per step (same seeds), OpenMP over problems, %d threads" % (nsample, cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * tot / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl.describe(nsample), "problems_per_step": nsample,
                   "note": "CPU port of the reference algorithm (oracle/, -O3, OpenMP over problems); the reference itself "
                           "needs Eigen+Boost(+Pinocchio), which this image lacks"},
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
    ap.add_argument("--workload", default="unicycle", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="problems per GPU (weak scaling)")
    ap.add_argument("--batch-total", type=int, default=0, help="total problems split over the ranks (strong scaling)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    wl = WORKLOADS[args.workload]()
    if args.impl == "reference":
        run_reference(args, wl, rank, world)
        return

    import ctypes as C

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
    from crocoddyl_b200 import sharding

    cb.set_device(local_rank)
    torch.cuda.set_device(local_rank)
    if args.batch_total:  # strong scaling: one global batch, contiguous shards
        total, scaling = args.batch_total, "strong"
        lo, hi = sharding.shard_range(total, rank, world)
        x0, xs_np, us_np = wl.inputs(lo, hi, total, 0)
    else:  # weak scaling: the same per-GPU batch on every rank, different seeds
        B0 = args.batch or wl.default_batch
        total, scaling = B0 * world, "weak"
        lo, hi = rank * B0, (rank + 1) * B0
        x0, xs_np, us_np = wl.inputs(0, B0, B0, 1000 * rank)
    B = hi - lo
    model = wl.model()
    problem = cb.ShootingProblem(x0, [model] * T, wl.terminal(model))
    if us_np is None:
        us_np = np.repeat(problem.quasiStatic(x0)[:, None, :], T, 1)
    # pinned host buffers (what a caller hands to the reference-facing API)
    xs_h = torch.from_numpy(np.ascontiguousarray(xs_np)).pin_memory()
    us_h = torch.from_numpy(np.ascontiguousarray(us_np)).pin_memory()
    xs_out = torch.empty_like(xs_h).pin_memory()
    us_out = torch.empty_like(us_h).pin_memory()
    solver = cb.SolverFDDP(problem)
    solver.setStorage(False)  # lean: keeps xs, us, K, k, Vx, Vxx, fs (the accessors north_star names)
    L = cb.lib()
    h = solver._h
    conv = np.zeros(B, dtype=np.uint8)
    scal = np.zeros((B, 17))
    dev = torch.device("cuda", local_rank)

    def dp(t):
        return C.cast(t.data_ptr(), C.POINTER(C.c_double))

    def gather_results():
        # the one collective of the path: per-problem {cost, iter, stop, converged} to every rank
        return sharding.gather_records(sharding.records_from_scalars(scal), total, dist, dev)

    def step_resident():
        cb.check(L.cgx_solver_solve_resident(h, 100, 0, wl.reginit))
        cb.check(L.cgx_solver_download(h, conv.ctypes.data_as(C.POINTER(C.c_ubyte))))
        cb.check(L.cgx_solver_get_scalars(h, cb.dptr(scal)))
        gather_results()
        return solver.timing()

    def step_e2e():
        # solve() + the xs/us getters as one reference-facing call (cgx_solver_solve_fetch): pinned host -> device ->
        # solve -> pinned host; the unicycle backend slices the batch so that the copies overlap the solve
        cb.check(L.cgx_solver_solve_fetch(h, dp(xs_h), dp(us_h), 100, 0, wl.reginit,
                                          conv.ctypes.data_as(C.POINTER(C.c_ubyte)), dp(xs_out), dp(us_out)))
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
    alg_bytes = float(T * (nb * wl.bytes_bwd + (nf - 1) * wl.bytes_fwd + wl.bytes_calc).sum())
    records = gather_results()
    # ---- e2e ---------------------------------------------------------------------------------
    for _ in range(2):
        step_e2e()
    sync()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    sync()
    e2e_ms = 1e3 * (time.perf_counter() - t0)
    # the fused call must have delivered exactly what the accessors hold (outside the timed region)
    chk = torch.empty_like(xs_out)
    cb.check(L.cgx_solver_get_all(h, b"xs", dp(chk)))
    if not torch.equal(chk, xs_out):
        raise RuntimeError("cgx_solver_solve_fetch: xs_out differs from cgx_solver_get_all(xs)")
    if dist is not None:
        v = torch.tensor([dev_ms, e2e_ms, wall_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        dev_ms, e2e_ms, wall_ms = [float(a) for a in v.cpu()]
    if rank == 0:
        solved = total * args.steps
        value = solved / (dev_ms * 1e-3)
        peak, peak_src = hbm_peak()
        kms = float(np.mean(kern_ms))
        achieved = alg_bytes / (kms * 1e-3) / 1e9
        h2d = int(xs_h.numel() * 8 + us_h.numel() * 8)
        d2h = int(xs_out.numel() * 8 + us_out.numel() * 8 + scal.size * 8 * 2 + B)
        ws_gb = (2 * xs_h.numel() + 2 * us_h.numel()) * 8 / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": wl.describe(B), "problems_per_gpu": B, "problems_total": total,
                       "storage": "lean (xs,us,K,k,Vx,Vxx,fs kept)",
                       "l2": "inputs larger than L2: trajectories alone are %.2f GB per GPU; every step rewrites the whole "
                             "solver state" % ws_gb,
                       "parallelism": "independent OCPs sharded over %d rank(s); NCCL all-gather of per-problem "
                       "{cost,iter,stop,converged}" % world},
            "e2e": {"value": solved / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / args.steps},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "wall_ms_per_step": wall_ms / args.steps,
            "solve_stats": {"iterations_mean": float(records[:, 1].mean()), "converged_frac": float(records[:, 3].mean()),
                            "backward_sweeps_mean": float(nb.mean()), "forward_sweeps_mean": float(nf.mean())},
            "roofline": {"kernel": wl.kernel, "bound": "hbm", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": kms},
        }
        if hasattr(wl, "flops_ric"):  # arm: the FP64 view of the same step (north_star: Riccati / ABA-derivative kernels)
            nrec = sg["n_records"]
            flops = float((T * nb * wl.flops_ric + (T + 1) * ((nrec + 1) * wl.flops_calcdiff + (nf + 1) * wl.flops_calc)).sum())
            fp64_peak = cb.probe_fp64_tflops()
            line["fp64"] = {"achieved": flops / (kms * 1e-3) / 1e12, "peak": fp64_peak, "unit": "TFLOP/s",
                            "frac": flops / (kms * 1e-3) / 1e12 / fp64_peak,
                            "peak_source": "measured on this GPU (cgx_probe_fp64_tflops, register-resident DFMA loop)",
                            "algorithmic_flops_per_step": flops,
                            "note": "upper bound of calcDiff evaluations: one per accepted iteration + the first"}
        if not args.no_cpu_baseline:
            nsample = wl.cpu_sample
            rate, cores, sec = cpu_oracle_rate(wl, nsample)
            line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "first %d problems of the same workload, %.2f s, OpenMP over problems" %
                                    (nsample, sec)}
        prof = os.path.join(ROOT, "profiles", "traffic_r01.json")
        if os.path.exists(prof) and wl.name == "unicycle" and B == 65536:
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
