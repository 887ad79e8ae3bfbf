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
    # algorithmic HBM bytes per node, SURVEY 8(d): B_fused = 8(n+m) read + 8(mn+3m+3n) written (K,k,Qu,Quuk,Vx,w,fs);
    # B_roll = 8(3n+3m+mn) per trial (read xs,us,K,k,fs ; write xs_try,us_try); first calc reads xs,us
    bytes_bwd = 8 * (3 + 2) + 8 * (2 * 3 + 3 * 2 + 3 * 3)  # 208
    bytes_fwd = 8 * (3 * 3 + 3 * 2 + 2 * 3)                # 168
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
    # (tools/count_fp64_ops.py: DFMA = 2, DADD = DMUL = 1 flop). The dense statement of the same algorithms run on a counting
    # scalar executes more (12,666 / 40,096: profiles/r02_arm_flops_counting_scalar.json); the smaller, issued figure is
    # the one used, so that the fraction is not inflated by work the kernel does not do.
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

    # NVML first (a thread polling every 5 ms: the timed region of the headline config is ~125 ms, in which
    # `nvidia-smi -lms 100` delivers one row), nvidia-smi as the fallback
    NVML_REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def _start_nvml(self):
        import pynvml
        pynvml.nvmlInit()
        self.nv = pynvml
        self.h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
        self.nv_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        self.nv_sm, self.nv_bits, self.nv_stop = [], 0, False

        def poll():
            get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            while not self.nv_stop:
                try:
                    self.nv_sm.append(float(pynvml.nvmlDeviceGetClockInfo(self.h, pynvml.NVML_CLOCK_SM)))
                    self.nv_bits |= int(get_reasons(self.h))
                except Exception:
                    pass
                time.sleep(0.005)

        self.th = threading.Thread(target=poll, daemon=True)
        self.th.start()

    def start(self):
        self.nv = None
        try:
            self._start_nvml()
            return
        except Exception:
            self.nv = None
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
        if getattr(self, "nv", None) is not None:
            self.nv_stop = True
            self.th.join(timeout=1.)
            reasons = sorted(n for n, bit in self.NVML_REASONS if self.nv_bits & bit)
            return {"sm_mhz": float(np.median(self.nv_sm)) if self.nv_sm else None, "sm_max_mhz": self.nv_max,
                    "reasons": reasons, "samples": len(self.nv_sm), "source": "nvml, 5 ms"}
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


def bind_to_gpu_numa_node(local_rank):
    """Before any pinned allocation: prefer host memory (and, where the cpuset allows it, CPUs) of the NUMA node the
    rank's GPU hangs off, so that the rank's H2D / D2H traffic does not cross the socket interconnect. Best effort:
    returns what was found / done for the JSON line."""
    import ctypes
    info = {"numa_node": None, "mempolicy": "default", "cpus_bound": False}
    try:
        import torch
        bdf = torch.cuda.get_device_properties(local_rank).pci_bus_id
        dom = getattr(torch.cuda.get_device_properties(local_rank), "pci_domain_id", 0)
        dev = getattr(torch.cuda.get_device_properties(local_rank), "pci_device_id", 0)
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bdf, dev)
        node = int(open(path).read().strip())
        info["numa_node"] = node
        if node < 0:
            return info
        libc = ctypes.CDLL(None, use_errno=True)
        mask = ctypes.c_ulong(1 << node)
        MPOL_PREFERRED, SYS_set_mempolicy = 1, 238  # x86-64
        rc = libc.syscall(SYS_set_mempolicy, MPOL_PREFERRED, ctypes.byref(mask), ctypes.c_ulong(64))
        info["mempolicy"] = "preferred:%d" % node if rc == 0 else "set_mempolicy failed (errno %d)" % ctypes.get_errno()
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if allowed:
            os.sched_setaffinity(0, allowed)
            info["cpus_bound"] = True
    except Exception as e:  # no sysfs / not permitted: run unbound
        info["error"] = str(e)[:80]
    return info


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


def cpu_oracle_rate(wl, nsample, nthreads=0, reps=1, want_scalars=False):
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
    if want_scalars:
        return nsample / best, cores, best, ob.scalars()
    return nsample / best, cores, best


def build_oracle_only():
    """The CPU legs need the oracle alone: make -C oracle (the product library is neither built nor loaded here)."""
    out = subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if out.returncode != 0:
        sys.stderr.write(out.stdout.decode())
        raise RuntimeError("oracle build failed")


def run_reference(args, wl, rank, world):
    """--impl reference: the reference algorithm's CPU implementation (oracle port, -O3, OpenMP over problems, all host
    threads) on the SAME workload as the GPU arm at N = 1: every step solves all default_batch problems of config C2
    (65,536; about 4 s per step on 16 threads). Rank 0 only; never touches the product library."""
    if rank != 0:
        return
    build_oracle_only()
    import oracle_lib as ol

    nsample = args.batch or wl.default_batch
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
    sample = "all %d problems of the workload per step (same seeds as the GPU arm), OpenMP over problems, %d threads" % (nsample, cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * tot / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl.describe(nsample), "problems_per_gpu": nsample, "problems_total": nsample,
                   "note": "CPU port of the reference algorithm (oracle/, -O3, OpenMP over problems, pinned to the reference "
                           "by tests/golden/ref_golden.npz); the reference itself needs Eigen+Boost(+Pinocchio), which this "
                           "image lacks. One host: the rate does not grow with --gpus."},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "sample_frac": 1.0},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


class Runner(object):
    """One workload on this rank's GPU: resident (`value`) and end-to-end timing of `steps` solves, max over ranks."""

    def __init__(self, wl, rank, local_rank, world, dist, batch=0, batch_total=0):
        import torch
        import crocoddyl_b200 as cb
        from crocoddyl_b200 import sharding

        self.wl, self.rank, self.world, self.dist = wl, rank, world, dist
        self.torch, self.cb, self.sharding = torch, cb, sharding
        if batch_total:  # strong scaling: one global batch, contiguous shards
            self.total, self.scaling = batch_total, "strong"
            lo, hi = sharding.shard_range(batch_total, rank, world)
            x0, xs_np, us_np = wl.inputs(lo, hi, batch_total, 0)
        else:  # weak scaling: the same per-GPU batch on every rank, different seeds
            B0 = batch or wl.default_batch
            self.total, self.scaling = B0 * world, "weak"
            x0, xs_np, us_np = wl.inputs(0, B0, B0, 1000 * rank)
        self.B = B = x0.shape[0]
        model = wl.model()
        self.problem = cb.ShootingProblem(x0, [model] * T, wl.terminal(model))
        if us_np is None:
            us_np = np.repeat(self.problem.quasiStatic(x0)[:, None, :], T, 1)
        # pinned host buffers (what a caller hands to the reference-facing API)
        self.xs_h = torch.from_numpy(np.ascontiguousarray(xs_np)).pin_memory()
        self.us_h = torch.from_numpy(np.ascontiguousarray(us_np)).pin_memory()
        self.xs_out = torch.empty_like(self.xs_h).pin_memory()
        self.us_out = torch.empty_like(self.us_h).pin_memory()
        self.solver = cb.SolverFDDP(self.problem)
        self.solver.setStorage(False)  # lean: keeps xs, us, K, k, Vx, Vxx, fs (the accessors north_star names)
        self.conv = np.zeros(B, dtype=np.uint8)
        self.scal = np.zeros((B, 17))
        self.dev = torch.device("cuda", local_rank)

    def dp(self, t):
        import ctypes as C
        return C.cast(t.data_ptr(), C.POINTER(C.c_double))

    def gather_results(self):
        # the one collective of the path: per-problem {cost, iter, stop, converged} to every rank
        return self.sharding.gather_records(self.sharding.records_from_scalars(self.scal), self.total, self.dist, self.dev)

    def step_resident(self):
        import ctypes as C
        cb, L, h = self.cb, self.cb.lib(), self.solver._h
        cb.check(L.cgx_solver_solve_resident(h, 100, 0, self.wl.reginit))
        cb.check(L.cgx_solver_download(h, self.conv.ctypes.data_as(C.POINTER(C.c_ubyte))))
        cb.check(L.cgx_solver_get_scalars(h, cb.dptr(self.scal)))
        self.gather_results()
        return self.solver.timing()

    def step_e2e(self):
        # solve() + the xs/us getters as one reference-facing call (cgx_solver_solve_fetch): pinned host -> device ->
        # solve -> pinned host; the unicycle backend slices the batch so that the copies overlap the solve
        import ctypes as C
        cb, L, h = self.cb, self.cb.lib(), self.solver._h
        cb.check(L.cgx_solver_solve_fetch(h, self.dp(self.xs_h), self.dp(self.us_h), 100, 0, self.wl.reginit,
                                          self.conv.ctypes.data_as(C.POINTER(C.c_ubyte)), self.dp(self.xs_out),
                                          self.dp(self.us_out)))
        cb.check(L.cgx_solver_get_scalars(h, cb.dptr(self.scal)))
        self.gather_results()

    def sync(self):
        self.torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def run(self, steps, warmup, sampler=None):
        torch, cb, wl, dist = self.torch, self.cb, self.wl, self.dist
        L, h = cb.lib(), self.solver._h
        cb.check(L.cgx_solver_upload(h, self.dp(self.xs_h), self.dp(self.us_h)))
        for _ in range(warmup):
            self.step_resident()
        if sampler is not None:
            sampler.start()
        self.sync()
        call_ms, kern_ms, launches = [], [], 0
        t0 = time.perf_counter()
        for _ in range(steps):
            tm = self.step_resident()
            call_ms.append(tm["call_ms"])
            kern_ms.append(tm["kernel_ms"])
            launches += tm["launches"] + 1  # + the scalar export kernel
        self.sync()
        wall_ms = 1e3 * (time.perf_counter() - t0)
        clocks = sampler.stop() if sampler is not None else None
        dev_ms = float(np.sum(call_ms))
        sg = self.solver.scalars()
        nb, nf = sg["n_backward"], sg["n_forward"]
        alg_bytes = float(T * (nb * wl.bytes_bwd + (nf - 1) * wl.bytes_fwd + wl.bytes_calc).sum())
        records = self.gather_results()
        for _ in range(max(2, min(warmup, 3))):
            self.step_e2e()
        self.sync()
        t0 = time.perf_counter()
        for _ in range(steps):
            self.step_e2e()
        self.sync()
        e2e_ms = 1e3 * (time.perf_counter() - t0)
        # the fused call must have delivered exactly what the accessors hold (outside the timed region)
        chk = torch.empty_like(self.xs_out)
        cb.check(L.cgx_solver_get_all(h, b"xs", self.dp(chk)))
        if not torch.equal(chk, self.xs_out):
            raise RuntimeError("cgx_solver_solve_fetch: xs_out differs from cgx_solver_get_all(xs)")
        # host-copy ceiling: the same pinned buffers moved H2D + D2H with nothing else running, all ranks at once
        d_x = torch.empty_like(self.xs_h, device=self.dev)
        d_u = torch.empty_like(self.us_h, device=self.dev)
        self.sync()
        t0 = time.perf_counter()
        for _ in range(3):
            d_x.copy_(self.xs_h, non_blocking=True)
            d_u.copy_(self.us_h, non_blocking=True)
            self.xs_out.copy_(d_x, non_blocking=True)
            self.us_out.copy_(d_u, non_blocking=True)
        self.sync()
        copy_ms = 1e3 * (time.perf_counter() - t0) / 3
        del d_x, d_u
        if dist is not None:
            v = torch.tensor([dev_ms, e2e_ms, wall_ms, copy_ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(v, op=dist.ReduceOp.MAX)
            dev_ms, e2e_ms, wall_ms, copy_ms = [float(a) for a in v.cpu()]
        solved = self.total * steps
        kms = float(np.mean(kern_ms))
        h2d = int(self.xs_h.numel() * 8 + self.us_h.numel() * 8)
        d2h = int(self.xs_out.numel() * 8 + self.us_out.numel() * 8 + self.scal.size * 8 * 2 + self.B)
        res = {
            "value": solved / (dev_ms * 1e-3), "ms_per_step": dev_ms / steps, "scaling": self.scaling,
            "problems_per_gpu": self.B, "problems_total": self.total,
            "e2e": {"value": solved / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / steps,
                    "host_copy_gbs_per_rank": (h2d + d2h) / (e2e_ms / steps * 1e-3) / 1e9,
                    "copy_only_ms_per_step": copy_ms,
                    "copy_only_gbs_per_rank": 2 * (self.xs_h.numel() + self.us_h.numel()) * 8 / (copy_ms * 1e-3) / 1e9},
            "gpu_launches": int(launches), "clocks": clocks, "wall_ms_per_step": wall_ms / steps,
            "solve_stats": {"iterations_mean": float(records[:, 1].mean()), "converged_frac": float(records[:, 3].mean()),
                            "backward_sweeps_mean": float(nb.mean()), "forward_sweeps_mean": float(nf.mean())},
            "kernel_ms": kms, "algorithmic_bytes_per_launch": alg_bytes, "scalars": sg,
        }
        return res


def fp64_view(wl, res, fp64_peak):
    sg = res["scalars"]
    nb, nf, nrec = sg["n_backward"], sg["n_forward"], sg["n_records"]
    flops = float((T * nb * wl.flops_ric + (T + 1) * ((nrec + 1) * wl.flops_calcdiff + (nf + 1) * wl.flops_calc)).sum())
    tf = flops / (res["kernel_ms"] * 1e-3) / 1e12
    return {"achieved": tf, "peak": fp64_peak, "unit": "TFLOP/s", "frac": tf / fp64_peak,
            "peak_source": "measured on this GPU (cgx_probe_fp64_tflops, register-resident DFMA loop)",
            "algorithmic_flops_per_step": flops,
            "note": "this rank's problems; upper bound of calcDiff evaluations: one per accepted iteration + the first"}


def traffic_record(kernel_key):
    """DRAM bytes per launch of the dominant kernel from the ncu --set full capture of the CURRENT kernels, if committed."""
    prof = os.path.join(ROOT, "profiles", "r02_traffic.json")
    if os.path.exists(prof):
        try:
            d = json.load(open(prof))
            if kernel_key in d:
                return d[kernel_key].get("dram_bytes_per_launch"), d[kernel_key].get("source")
        except Exception:
            pass
    return None, None


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
    ap.add_argument("--no-extras", action="store_true", help="only the headline workload (no C1 / C3 / C4 / C5 blocks)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    wl = WORKLOADS[args.workload]()
    if args.impl == "reference":
        run_reference(args, wl, rank, world)
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
    numa = bind_to_gpu_numa_node(local_rank)
    # ---- headline: BASELINE config C2 (or --workload arm) ---------------------------------------------------------
    runner = Runner(wl, rank, local_rank, world, dist, args.batch, args.batch_total)
    B, total = runner.B, runner.total
    sampler = ClockSampler(local_rank) if rank == 0 else None
    res = runner.run(args.steps, args.warmup, sampler)
    line = None
    if rank == 0:
        peak, peak_src = hbm_peak()
        achieved = res["algorithmic_bytes_per_launch"] / (res["kernel_ms"] * 1e-3) / 1e9
        ws_gb = (2 * runner.xs_h.numel() + 2 * runner.us_h.numel()) * 8 / 1e9
        traffic, traffic_src = traffic_record(wl.name) if (B == wl.default_batch) else (None, None)
        line = {
            "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": res["scaling"],
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": wl.describe(B), "problems_per_gpu": B, "problems_total": total,
                       "storage": "lean (xs,us,K,k,Vx,Vxx,fs kept)",
                       "l2": "inputs larger than L2: trajectories alone are %.2f GB per GPU; every step rewrites the whole "
                             "solver state" % ws_gb,
                       "parallelism": "independent OCPs sharded over %d rank(s); NCCL all-gather of per-problem "
                       "{cost,iter,stop,converged}" % world},
            "e2e": res["e2e"], "host_binding_rank0": numa, "gpu_launches": res["gpu_launches"], "clocks": res["clocks"],
            "wall_ms_per_step": res["wall_ms_per_step"], "solve_stats": res["solve_stats"],
            "roofline": {"kernel": wl.kernel, "bound": "hbm", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": res["algorithmic_bytes_per_launch"],
                         "bytes_per_node": {"backward": wl.bytes_bwd, "forward": wl.bytes_fwd, "first_calc": wl.bytes_calc,
                                            "source": "SURVEY 8(d): B_fused / B_roll"},
                         "kernel_ms": res["kernel_ms"]},
        }
        fp64_peak = cb.probe_fp64_tflops()
        if hasattr(wl, "flops_ric"):
            line["fp64"] = fp64_view(wl, res, fp64_peak)
        if not args.no_cpu_baseline:
            nsample = min(wl.cpu_sample, B)
            rate, cores, sec, so = cpu_oracle_rate(wl, nsample, want_scalars=True)
            line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "first %d problems of the same workload, %.2f s, OpenMP over problems" %
                                    (nsample, sec), "sample_frac": nsample / float(B)}
            if not args.batch_total:  # the sample is the head of rank 0's own problems: how often do the counts agree?
                same = so["n_records"] == res["scalars"]["n_records"][:nsample]
                line["solve_stats"]["iter_parity_frac"] = float(same.mean())
                line["solve_stats"]["iter_parity_note"] = ("identical iteration count GPU vs CPU oracle on the cpu_baseline "
                                                           "sample; floor set by libm rounding: profiles/r02_sincos_sensitivity.json")
    # ---- the other BASELINE configs, same process, same JSON line -------------------------------------------------
    if not args.no_extras and args.workload == "unicycle" and not args.batch and not args.batch_total:
        del runner
        extras = run_extras(args, rank, local_rank, world, dist, cb, line)
        if rank == 0:
            line.update(extras)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def run_extras(args, rank, local_rank, world, dist, cb, line):
    """C1 (single LQR OCP, ms per call), C3 (Riccati us per problem-sweep, FP64 fraction), C4 (arm B = 1024) at N = 1;
    C5 (arm, B = 8192 SPLIT over the ranks: strong scaling) at every N."""
    out = {}
    arm = Arm()
    steps, warmup = min(args.steps, 3), 3
    fp64_peak = cb.probe_fp64_tflops() if rank == 0 else None
    if world == 1:
        import bench_riccati

        ric = bench_riccati.run(peak_tflops=fp64_peak)
        out["riccati"] = [{"nx": r["nx"], "nu": r["nu"], "B": r["B"], "us_per_iter": r["us_per_problem_sweep"],
                           "sweep_ms_batched": r["sweep_ms_batched"], "fp64_tflops": r["fp64_tflops"], "frac": r["frac"]}
                          for r in ric]
        out["riccati_note"] = ("C3: one SolverDDP::backwardPass per problem, LQR nx in {12,36,72}, nu = nx/3, T=100, B=4096; "
                               "us_per_iter = batched sweep time / B; frac = F_ric flops / measured DFMA peak %.1f TFLOP/s" % fp64_peak)
        c1 = bench_riccati.run_c1(trials=300, cpu_trials=100)
        out["c1"] = {"ms_per_call": c1["gpu_call"]["mean_ms"], "ms_min": c1["gpu_call"]["min_ms"], "ms_max": c1["gpu_call"]["max_ms"],
                     "kernel_ms": c1["gpu_kernel"]["mean_ms"], "cpu_port_1thread_ms": c1["cpu_oracle_1thread"]["mean_ms"],
                     "config": c1["config"]}
        r4 = Runner(arm, rank, local_rank, world, dist, batch=1024)
        a4 = r4.run(steps, warmup)
        out["arm_c4"] = {"solves_per_s": a4["value"], "ms_per_step": a4["ms_per_step"], "e2e": a4["e2e"],
                         "fp64": fp64_view(arm, a4, fp64_peak), "solve_stats": a4["solve_stats"],
                         "gpu_launches": a4["gpu_launches"], "config": arm.describe(1024)}
        if not args.no_cpu_baseline:
            rate, cores, sec, _ = cpu_oracle_rate(arm, arm.cpu_sample, want_scalars=True)
            out["arm_c4"]["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                                             "sample": "first %d problems, %.2f s" % (arm.cpu_sample, sec)}
        del r4
    r5 = Runner(arm, rank, local_rank, world, dist, batch_total=8192)
    a5 = r5.run(steps, warmup)
    if rank == 0:
        out["arm_c5"] = {"solves_per_s": a5["value"], "ms_per_step": a5["ms_per_step"], "e2e": a5["e2e"], "scaling": "strong",
                         "problems_total": 8192, "problems_per_gpu": a5["problems_per_gpu"], "n_gpus": world,
                         "fp64": fp64_view(arm, a5, fp64_peak), "solve_stats": a5["solve_stats"],
                         "strong_efficiency_inputs": "solves_per_s at n_gpus = N over N x solves_per_s at n_gpus = 1",
                         "config": "C5: " + arm.describe(a5["problems_per_gpu"]) + "; 8192 problems split contiguously over the ranks"}
    return out


if __name__ == "__main__":
    main()
