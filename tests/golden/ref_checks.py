"""Checks of an implementation of the solve path against tests/golden/ref_golden.npz — vectors produced by the
reference's own Python statements of the path (tests/golden/make_ref_golden.py: the legacy NumPy crocoddyl under
unittest/python/crocoddyl and DDPDerived/FDDPDerived of bindings/python/crocoddyl/utils).

The same checks run against the CPU oracle (tests/test_ref_golden_cpu.py, `OracleEngine`) and against the CUDA path
through the C ABI (tests/test_ref_golden_gpu.py, `GpuEngine`), so both are pinned to the reference and not to each
other. Tolerance: 1e-9 relative (BASELINE.json north_star), discrete decisions exact.
"""
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_golden.npz")
RTOL = 1e-9
_cache = {}


def load():
    if "g" not in _cache:
        z = np.load(GOLD)
        cases = {}
        for key in z.files:
            c, k = key.split("/", 1)
            cases.setdefault(c, {})[k] = z[key]
        _cache["g"] = cases
    return _cache["g"]


def names(prefix):
    return sorted(c for c in load() if c.startswith(prefix))


def close(a, b, what, rtol=RTOL):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    scale = max(1.0, float(np.max(np.abs(b)))) if b.size else 1.0
    np.testing.assert_allclose(a, b, rtol=rtol, atol=rtol * scale, err_msg=what)


def lqr_pack(g):
    """C-ABI / oracle parameter block of an LQR model from the matrices stored in a golden case."""
    f = lambda a: np.asarray(a, dtype=np.float64).ravel(order="F")
    return np.concatenate([[float(g["drift_free"])], f(g["p_Fx"]), f(g["p_Fu"]), f(g["p_f0"]), f(g["p_Lxx"]),
                           f(g["p_Lxu"]), f(g["p_Luu"]), f(g["p_lx"]), f(g["p_lu"])])


UNICYCLE_PARAMS = np.array([0.1, 10.0, 1.0])  # dt, w_x, w_u (unicycle.hxx:14-15)


def model_of(cname, g):
    """-> (kind, params, n, m)"""
    if "unicycle" in cname:
        return "unicycle", UNICYCLE_PARAMS, 3, 2
    n, m = g["p_Fu"].shape
    return "lqr", lqr_pack(g), n, m


# ------------------------------------------------------------------------------------------------------------------
# engines: .create(kind, solver, T, n, m, params, x0[B,n]) -> object with the methods used below
# ------------------------------------------------------------------------------------------------------------------
class OracleEngine:
    def __init__(self, kind, solver, T, n, m, params, x0):
        import oracle_lib as ol
        self.ob = ol.OracleBatch(kind, solver, T, n, m, params, np.atleast_2d(x0))
        self.T = T

    def set_candidate(self, xs, us, feasible):
        self.ob.set_candidate(xs, us, feasible)

    def set_reg(self, xreg, ureg):
        self.ob.set("xreg", xreg)
        self.ob.set("ureg", ureg)

    def problem_calc(self, xs, us):
        self.ob.set_candidate(xs, us, False)
        return self.ob.calc()

    def problem_calcdiff(self, xs, us):  # after problem_calc
        return self.ob.calcdiff()

    def node(self, field, t):
        if field == "cost":
            return self.ob.get("node_cost", 0, t)[0]
        return self.ob.get(field, 0, t)

    def compute_direction(self):
        self.ob.calcdiff()
        assert not self.ob.backward().any()

    def get(self, name):
        return self.ob.get_all(name)

    def try_step(self, a):
        dV, st = self.ob.try_step(a)
        assert not st.any()
        return dV

    def expected_improvement(self):
        return self.ob.expected_improvement()[0]

    def stopping_criteria(self):
        return self.ob.expected_improvement()[1]

    def cost_try(self):
        return self.ob.scalars()["cost_try"]

    def cost(self):
        return self.ob.scalars()["cost"]

    def solve(self, xs, us, maxiter, feasible, reginit):
        self.ob.solve(xs, us, maxiter, feasible, reginit)
        return self.ob.scalars()

    def trace(self, b):
        return self.ob.trace(b)

    def rollout(self, us):
        return None  # the oracle's C API has no separate rollout entry; covered by try_step on a feasible start


class GpuEngine:
    def __init__(self, kind, solver, T, n, m, params, x0):
        import crocoddyl_b200 as cb
        x0 = np.atleast_2d(x0)
        if kind == "unicycle":
            mod = cb.ActionModelUnicycle()
        else:
            mod = _RawLqr(n, m, params)
        self.problem = cb.ShootingProblem(x0, [mod] * T, mod)
        self.solver = (cb.SolverFDDP if solver == "fddp" else cb.SolverDDP)(self.problem)
        self.T = T

    def set_candidate(self, xs, us, feasible):
        self.solver.setCandidate(xs, us, feasible)

    def set_reg(self, xreg, ureg):
        self.solver.x_reg = xreg
        self.solver.u_reg = ureg

    def problem_calc(self, xs, us):
        self._xu = (xs, us)
        return self.problem.calc(xs, us)

    def problem_calcdiff(self, xs, us):
        return self.problem.calcDiff(xs, us)

    def node(self, field, t):
        v = self.problem.data(field, 0, t)
        return v[0] if field == "cost" else v

    def compute_direction(self):
        self.solver.computeDirection(True)

    def get(self, name):
        return self.solver.get(name)

    def try_step(self, a):
        return self.solver.tryStep(a)

    def expected_improvement(self):
        return self.solver.expectedImprovement()

    def stopping_criteria(self):
        return self.solver.stoppingCriteria()

    def cost_try(self):
        return self.solver.scalars()["cost_try"]

    def cost(self):
        return self.solver.scalars()["cost"]

    def solve(self, xs, us, maxiter, feasible, reginit):
        self.solver.enableTrace(maxiter + 1)
        self.solver.solve(xs, us, maxiter, feasible, reginit)
        return self.solver.scalars()

    def trace(self, b):
        return self.solver.trace(b)

    def rollout(self, us):
        return self.problem.rollout(us)


class _RawLqr:
    """LQR node model given directly by its C-ABI parameter block ([nparams] or [B, nparams])."""
    kind = 0

    def __init__(self, n, m, params):
        self.nx = self.ndx = n
        self.nu = m
        self.nr = 0
        self._p = np.ascontiguousarray(params, dtype=np.float64)

    def params(self):
        return self._p


# ------------------------------------------------------------------------------------------------------------------
# checks
# ------------------------------------------------------------------------------------------------------------------
def check_node(Engine, cname):
    """ActionModelUnicycle / ActionModelLQR calc + calcDiff per node (legacy unicycle.py:24-50, action.py:140-172)."""
    g = load()[cname]
    kind, params, n, m = model_of(cname, g)
    X, U = g["x"], g["u"]
    N = X.shape[0]
    xs = np.concatenate([X, X[:1]], 0)[None]  # node t evaluates (X[t], U[t]); terminal node evaluates X[0]
    e = Engine(kind, "ddp", N, n, m, params, X[0])
    e.problem_calc(xs, U[None])
    e.problem_calcdiff(xs, U[None])
    for t in range(N):
        close(e.node("cost", t), g["cost"][t], "%s cost[%d]" % (cname, t))
        for f in ("xnext", "Lx", "Lu") + (("Fx", "Fu", "Lxx", "Lxu", "Luu") if kind == "unicycle" else ()):
            close(e.node(f, t), g[f][t], "%s %s[%d]" % (cname, f, t))
        if kind == "lqr":  # the data copies of the model matrices (lqr.hxx:65-69)
            for f in ("Fx", "Fu", "Lxx", "Lxu", "Luu"):
                close(e.node(f, t), g["p_" + f], "%s %s[%d]" % (cname, f, t))
    if kind == "unicycle":  # terminal call with u = unone
        close(e.node("cost", N), g["term_cost"], cname + " terminal cost")
        close(e.node("Lx", N), g["term_Lx"], cname + " terminal Lx")
        close(e.node("Lxx", N), g["term_Lxx"], cname + " terminal Lxx")


def check_step(Engine, cname):
    """ShootingProblem calc/calcDiff/rollout, SolverDDP/FDDP computeDirection (calcDiff + gaps + backwardPass),
    tryStep (forwardPass) and expectedImprovement at one candidate (legacy shooting.py, ddp.py:40-83,184-274,
    fddp.py:50-124,236-338)."""
    g = load()[cname]
    kind, params, n, m = model_of(cname, g)
    T = g["us"].shape[0]
    xs, us = g["xs"][None], g["us"][None]
    feasible = bool(g["feasible"])
    for sname in ("ddp", "fddp"):
        e = Engine(kind, sname, T, n, m, params, g["x0"])
        if sname == "ddp":
            close(e.problem_calc(xs, us), [g["calc_cost"]], cname + " problem.calc")
            close(e.problem_calcdiff(xs, us), [g["calcdiff_cost"]], cname + " problem.calcDiff")
            for f in ("xnext", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu"):
                for t in range(T):
                    close(e.node(f, t), g["d_" + f][t], "%s data[%d].%s" % (cname, t, f))
            close(e.node("Lx", T), g["dT_Lx"], cname + " terminal Lx")
            close(e.node("Lxx", T), g["dT_Lxx"], cname + " terminal Lxx")
            r = e.rollout(us)
            if r is not None:
                close(r[0], g["rollout"], cname + " rollout")
        e.set_candidate(xs, us, feasible)
        e.set_reg(float(g["xreg"]), float(g["ureg"]))
        e.compute_direction()
        close(e.cost(), [g[sname + "_cost"]], cname + " cost")
        fs = e.get("fs")[0]
        close(fs, g[sname + "_gaps"], "%s %s gaps" % (cname, sname))
        for f in ("Vxx", "Qxx", "Qxu", "Quu", "Qx", "Qu", "K", "k"):
            close(e.get(f)[0], g[sname + "_" + f], "%s %s %s" % (cname, sname, f))
        Vx, Vxx = e.get("Vx")[0], e.get("Vxx")[0]
        if sname == "ddp" and not feasible:
            # the legacy DDP keeps Vx at the candidate and adds Vxx*gap inside Qx/Qu (ddp.py:205-208); the C++ (and the
            # legacy FDDP) store Vx + Vxx*gap (ddp.cpp:189-191,242-244)
            Vx = Vx - np.einsum("tij,tj->ti", Vxx, fs)
        close(Vx, g[sname + "_Vx"], "%s %s Vx" % (cname, sname))
        close(e.stopping_criteria(), [g[sname + "_stop"]], "%s %s stop" % (cname, sname))
        if sname == "ddp":
            close(e.expected_improvement()[0], g["ddp_d"], cname + " ddp expectedImprovement")
        for a in (1.0, 0.5, 0.125):
            tag = "%s_a%g_" % (sname, a)
            dV = e.try_step(a)
            close(e.get("xs_try")[0], g[tag + "xs_try"], cname + " " + tag + "xs_try")
            close(e.get("us_try")[0], g[tag + "us_try"], cname + " " + tag + "us_try")
            close(e.cost_try(), [g[tag + "cost_try"]], cname + " " + tag + "cost_try")
            # dV = cost - cost_try: absolute tolerance on the scale of the costs
            np.testing.assert_allclose(dV[0], g[tag + "dV"], rtol=RTOL, atol=RTOL * abs(float(g[sname + "_cost"])))
            if sname == "fddp":
                d = e.expected_improvement()[0]
                np.testing.assert_allclose(d, g[tag + "d"], rtol=1e-8, atol=RTOL * abs(float(g[sname + "_cost"])))


def check_kkt(Engine, cname):
    """DDP on an LQR problem reaches the SolverKKT solution (legacy kkt.py; unittest/test_solvers.cpp:65-110)."""
    g = load()[cname]
    kind, params, n, m = model_of(cname, g)
    T = g["us"].shape[0]
    x0 = g["x0"]
    xs0, us0 = np.repeat(x0[None, None, :], T + 1, 1), np.zeros((1, T, m))
    for sname in ("ddp", "fddp"):
        e = Engine(kind, sname, T, n, m, params, x0)
        sc = e.solve(xs0, us0, 10, False, 1e-9)
        assert sc["converged"][0] == 1
        close(e.get("xs")[0], g["xs"], cname + " xs vs KKT", 1e-8)
        close(e.get("us")[0], g["us"], cname + " us vs KKT", 1e-8)
        close(sc["cost"], [g["cost"]], cname + " cost vs KKT", 1e-9)


def check_solve(Engine, cname, traj_rtol=RTOL):
    """Whole solve() against DDPDerived / FDDPDerived (bindings/python/crocoddyl/utils/__init__.py:842-1187): same
    number of iterations, same accepted step length, regularisation and feasibility in every iteration; cost per
    iteration and the final xs, us, K, k, Vx, Vxx within tolerance."""
    g = load()[cname]
    kind, params, n, m = model_of(cname, g)
    sname = "fddp" if "_fddp_" in cname else "ddp"
    T = g["init_us"].shape[0]
    e = Engine(kind, sname, T, n, m, params, g["x0"])
    maxiter = 100 if kind == "unicycle" else 10
    sc = e.solve(g["init_xs"][None], g["init_us"][None], maxiter, False, 1e-9)
    ref = g["trace"]  # iter cost stop x_reg u_reg stepLength isFeasible dV dV_exp
    tr = e.trace(0)   # iter cost stop d0 d1 xreg ureg steplength is_feasible dV dVexp alpha_idx retries
    assert tr.shape[0] == ref.shape[0], "%s: %d iterations, the reference took %d" % (cname, tr.shape[0], ref.shape[0])
    assert sc["converged"][0] == g["converged"]
    for col_t, col_r, what in ((0, 0, "iter"), (5, 3, "xreg"), (6, 4, "ureg"), (7, 5, "steplength"), (8, 6, "feasible")):
        assert np.array_equal(tr[:, col_t], ref[:, col_r]), (cname, what, tr[:, col_t], ref[:, col_r])
    np.testing.assert_allclose(tr[:, 1], ref[:, 1], rtol=RTOL, err_msg=cname + " cost per iteration")
    close(sc["cost"], [g["cost"]], cname + " cost")
    for f in ("xs", "us"):
        close(e.get(f)[0], g[f], "%s %s" % (cname, f), traj_rtol)
    # gains / value function of the LAST backward pass (taken at the candidate before the final step)
    for f in ("K", "k", "Vxx", "Vx", "Qu", "Quu", "Qxx", "Qxu", "Qx"):
        close(e.get(f)[0], g[f], "%s %s" % (cname, f), max(traj_rtol, 1e-8) if f in ("k", "Vx", "Qu", "Qx") else traj_rtol)
    return tr, ref
