/*
 * crocoddyl_b200 — C ABI of the B200-native batched DDP/FDDP solve path.
 *
 * The reference (julesser/crocoddyl, C++/Boost.Python) has no FFI for this path: its boundary is the
 * C++ class interface ShootingProblem / SolverDDP / SolverFDDP. Every entry point below names the
 * reference member it replaces (paths relative to the reference tree). A "problem" here is a BATCH of
 * B independent ShootingProblems that share T, the node-model family and (optionally) its parameters;
 * B = 1 behaves like the reference classes. The C++ facade with the reference's own class names is
 * include/crocoddyl_b200.hpp; INTEGRATION.md shows the reference-side binding.
 *
 * Conventions
 *   - all numbers are FP64; vectors are contiguous; matrices are COLUMN-MAJOR (Eigen::MatrixXd default,
 *     include/crocoddyl/core/mathbase.hpp), K is nu x ndx, Qxu/Fu/Lxu are ndx x nu;
 *   - host trajectories: xs = [B][T+1][nx], us = [B][T][nu] (std::vector<Eigen::VectorXd> per problem,
 *     problems concatenated);
 *   - every function returns CGX_OK (0) or a negative error code and stores a message retrievable with
 *     cgx_last_error() (mirrors crocoddyl::Exception text, include/crocoddyl/core/utils/exception.hpp:23-29);
 *     no C++ exception crosses this boundary;
 *   - one host thread per handle; each solver owns one CUDA stream (reference: not thread-safe either);
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails with CGX_ERR_CUDA.
 */
#ifndef CROCODDYL_B200_H_
#define CROCODDYL_B200_H_

#include <stddef.h>

#if defined(__GNUC__)
#define CGX_API __attribute__((visibility("default")))
#else
#define CGX_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cgx_problem_s* cgx_problem_t;
typedef struct cgx_solver_s* cgx_solver_t;

enum cgx_status {
  CGX_OK = 0,
  CGX_ERR_INVALID = -1, /* invalid argument / dimension (reference: throw_pretty "Invalid argument: ...") */
  CGX_ERR_CUDA = -2,    /* CUDA runtime failure or no device */
  CGX_ERR_UNSUPPORTED = -3,
  CGX_ERR_NOTFOUND = -4
};

/* node-model family of the batch */
enum cgx_model_kind {
  CGX_MODEL_LQR = 0,      /* ActionModelLQR       include/crocoddyl/core/actions/lqr.hxx:14-70 */
  CGX_MODEL_UNICYCLE = 1, /* ActionModelUnicycle  include/crocoddyl/core/actions/unicycle.hxx:13-73 */
  CGX_MODEL_ARM = 2       /* IntegratedActionModelEuler(DifferentialActionModelFreeFwdDynamics) on a fixed-base
                             revolute chain: include/crocoddyl/core/integrator/euler.hxx:41-131,
                             include/crocoddyl/multibody/actions/free-fwddyn.hxx:44-169 */
};
enum cgx_solver_kind {
  CGX_SOLVER_DDP = 0,     /* src/core/solvers/ddp.cpp */
  CGX_SOLVER_FDDP = 1,    /* src/core/solvers/fddp.cpp */
  CGX_SOLVER_BOXDDP = 2,  /* src/core/solvers/box-ddp.cpp  (control-limited, BoxQP in computeGains, clamp in forwardPass) */
  CGX_SOLVER_BOXFDDP = 3  /* src/core/solvers/box-fddp.cpp */
};

CGX_API const char* cgx_last_error(void);
/* number of visible CUDA devices (0 if none / no driver) and selection of the device used by handles
 * created afterwards by this host thread */
CGX_API int cgx_device_count(void);
CGX_API int cgx_set_device(int device);
CGX_API const char* cgx_version(void);

/* ----------------------------------------------------------------------------------------------
 * ShootingProblem  (include/crocoddyl/core/optctrl/shooting.hpp:48-63, shooting.hxx:22-68)
 *   kind/params describe the node model aliased into all T running nodes and the terminal node
 *   (benchmark/unicycle-optctrl.cpp:25-32, benchmark/lqr-optctrl.cpp:27-35, benchmark/factory/arm.hpp:60-90).
 *   params layout (doubles):
 *     LQR      [drift_free, Fx(n*n), Fu(n*m), f0(n), Lxx(n*n), Lxu(n*m), Luu(m*m), lx(n), lu(m)]
 *     UNICYCLE [dt, w_x, w_u]                                   (unicycle.hxx:14-15 defaults 0.1, 10, 1)
 *     ARM      [dt, gravity(3), w_gripperPose, w_xReg, w_uReg, frame parent joint, frame placement R(9 row-major) p(3),
 *               Mref R(9) p(3), goal cost kind, 3 reserved] + per joint [axis(3), placement in parent R(9) p(3), mass,
 *               COM lever(3), inertia about the COM Ixx Ixy Ixz Iyy Iyz Izz, armature, state-cost weight of q_i, of v_i]
 *               = cgx_arm_param_count(nv) = 36 + 28 nv doubles
 *               (the model of benchmark/factory/arm.hpp:60-90; the terminal node is the same model with dt = 0).
 *               goal cost kind: 0 CostModelFramePlacement (frame-placement.hxx), 1 CostModelFrameTranslation
 *               (frame-translation.hxx: only Mref p is used), 2 CostModelFrameRotation (frame-rotation.hxx: only Mref R).
 *               armature: free-fwddyn.hxx:66-74,108-114 (0 = none). State-cost weights: ActivationModelWeightedQuad
 *               (weighted-quadratic.hpp) on CostModelState; 1 = the plain ActivationModelQuad.
 *   per_problem_params != 0: params holds B consecutive blocks of nparams doubles, else one shared block.
 *   x0s = [B][nx].
 */
CGX_API int cgx_problem_create(cgx_problem_t* out, int kind, int B, int T, int nx, int nu, const double* params,
                       long nparams, int per_problem_params, const double* x0s);
CGX_API int cgx_problem_destroy(cgx_problem_t p);
/* A terminal model that differs from the running model: ShootingProblem(x0, running_models, terminal_model)
 * (shooting.hpp:48-63) with terminal_model != running_models[0], as examples/arm_manipulation.py:41-63 (terminal cost =
 * gripperPose only). tparams has the layout and size of the running block (same kind, nx; per_problem_params as given at
 * creation). Without this call the terminal node uses the running block — for CGX_MODEL_ARM with dt = 0, the model of
 * benchmark/factory/arm.hpp:89-90. The terminal block's dt is used as given (Euler with dt = 0 for a terminal knot). */
CGX_API int cgx_problem_set_terminal_params(cgx_problem_t p, const double* tparams, long nparams);
CGX_API int cgx_problem_get_terminal_params(cgx_problem_t p, double* tparams);
/* ActionModelAbstract::set_u_lb / set_u_ub of the running model (action-base.hxx:117-140): lb, ub = [nu] shared by the
 * batch (per_problem == 0) or [B][nu]; +-INFINITY = unbounded. has_control_limits follows the reference rule (some finite
 * entry in lb AND some finite entry in ub). Read by CGX_SOLVER_BOXDDP / CGX_SOLVER_BOXFDDP. */
CGX_API int cgx_problem_set_u_bounds(cgx_problem_t p, const double* lb, const double* ub, int per_problem);
CGX_API int cgx_problem_get_u_bounds(cgx_problem_t p, double* lb, double* ub, int* per_problem, int* has_control_limits);
CGX_API int cgx_problem_set_x0(cgx_problem_t p, const double* x0s); /* ShootingProblem::set_x0 shooting.hxx:393-400 */
CGX_API int cgx_problem_get_x0(cgx_problem_t p, double* x0s);       /* get_x0 shooting.hxx:325-328 */
/* get_T / get_nx / get_ndx / get_nu_max (shooting.hxx:319-322, 367-390) and the batch size */
CGX_API int cgx_problem_dims(cgx_problem_t p, int* B, int* T, int* nx, int* ndx, int* nu);
/* ShootingProblem::calc (shooting.hxx:133-161): fills every node's xnext/cost, costs[b] = total cost */
CGX_API int cgx_problem_calc(cgx_problem_t p, const double* xs, const double* us, double* costs);
/* ShootingProblem::calcDiff (shooting.hxx:164-195): fills Fx,Fu,Lx,Lu,Lxx,Lxu,Luu of every node;
 * costs[b] = sum of the node costs left by the last calc */
CGX_API int cgx_problem_calc_diff(cgx_problem_t p, const double* xs, const double* us, double* costs);
/* ShootingProblem::rollout (shooting.hxx:198-223): xs[b][0] = x0, xs[b][t+1] = xnext(xs[t], us[t]) */
CGX_API int cgx_problem_rollout(cgx_problem_t p, const double* us, double* xs);
/* ActionModelAbstract::quasiStatic for every problem at x (action-base.hxx:39-83 / free-fwddyn.hxx:146-169);
 * xs_point = [B][nx], us_out = [B][nu] */
CGX_API int cgx_problem_quasi_static(cgx_problem_t p, const double* xs_point, double* us_out);
/* ActionDataAbstract fields of node t of problem b after calc/calcDiff (core/action-base.hpp:101-142):
 * field in {"cost","xnext","Fx","Fu","Lx","Lu","Lxx","Lxu","Luu"}; t == T is the terminal data. "r" (the cost residual,
 * nr doubles) exists where the reference's model writes it: CGX_MODEL_UNICYCLE (nr = 5, unicycle.hxx:39-40), after calc.
 * Returns the number of doubles written (>0) or an error (<0). */
CGX_API int cgx_problem_get_data(cgx_problem_t p, const char* field, int b, int t, double* out);
/* ARM parameter block size for an nv-joint chain */
CGX_API long cgx_arm_param_count(int nv);

/* ----------------------------------------------------------------------------------------------
 * SolverDDP / SolverFDDP (include/crocoddyl/core/solvers/ddp.hpp:48-302, fddp.hpp:50-100)
 */
CGX_API int cgx_solver_create(cgx_solver_t* out, cgx_problem_t p, int solver_kind);
CGX_API int cgx_solver_destroy(cgx_solver_t s);

/* setters/getters of the scalar parameters: name in
 *   th_acceptstep th_stop (solver-base.hpp:217-257) th_grad th_stepdec th_stepinc regfactor regmin regmax
 *   (ddp.hpp:156-271) th_acceptnegstep (fddp.hpp:87-92) xreg ureg (solver-base.hpp:242-247)
 * Invalid values are rejected like the reference setters (ddp.cpp:420-486). */
CGX_API int cgx_solver_set_param(cgx_solver_t s, const char* name, double value);
CGX_API int cgx_solver_get_param(cgx_solver_t s, const char* name, double* value);
/* set_alphas / get_alphas (ddp.cpp:448-462); at most 16 step lengths */
CGX_API int cgx_solver_set_alphas(cgx_solver_t s, const double* alphas, int n);
CGX_API int cgx_solver_get_alphas(cgx_solver_t s, double* alphas, int* n);
/* Storage policy of the backward pass. full != 0 (default): Qxx,Qxu,Quu,Qx,Qu are kept per node and
 * readable afterwards (ddp.hpp:201-221). full == 0: only xs,us,K,k,Vx,Vxx,fs are kept (the accessors
 * BASELINE.json north_star names); the others return CGX_ERR_UNSUPPORTED. */
CGX_API int cgx_solver_set_storage(cgx_solver_t s, int full);
/* per-iteration records (what CallbackVerbose prints, src/core/utils/callbacks.cpp:37-46) kept on device for
 * every problem: on != 0 enables recording of up to max_records iterations per problem */
CGX_API int cgx_solver_enable_trace(cgx_solver_t s, int on, int max_records);

/* bool SolverDDP::solve / SolverFDDP::solve(init_xs, init_us, maxiter, is_feasible, reginit)
 * (ddp.cpp:41-118, fddp.cpp:19-105). init_xs == NULL: every state = zero (solver-base.cpp:46-50);
 * init_us == NULL: zeros. converged[b] (may be NULL) receives the bool the reference returns.
 * Host buffers; the call uploads, solves on the device and downloads the per-problem scalars. */
CGX_API int cgx_solver_solve(cgx_solver_t s, const double* init_xs, const double* init_us, int maxiter, int is_feasible,
                     double reginit, unsigned char* converged);
/* solve() followed by get_xs() / get_us() (solver-base.hpp:162-167) as ONE call: xs_out = [B][T+1][nx] and
 * us_out = [B][T][nu] (each may be NULL) receive the solver's xs / us after the solve. Same results as
 * cgx_solver_solve + cgx_solver_get_all; on the thread-per-problem backend the batch is processed in slices on
 * separate streams so that the host<->device copies of one slice overlap the solve of the others. */
CGX_API int cgx_solver_solve_fetch(cgx_solver_t s, const double* init_xs, const double* init_us, int maxiter,
                                   int is_feasible, double reginit, unsigned char* converged, double* xs_out,
                                   double* us_out);
/* The same solve split at the PCIe boundary (for pipelines that keep data resident in HBM):
 *   upload:          copies init_xs/init_us (host) into device staging buffers
 *   solve_resident:  setCandidate from the staging buffers + the whole solve loop, no host<->device copies;
 *                    returns after the kernels completed
 *   download:        converged flags (and the scalar block) back to the host */
CGX_API int cgx_solver_upload(cgx_solver_t s, const double* init_xs, const double* init_us);
CGX_API int cgx_solver_solve_resident(cgx_solver_t s, int maxiter, int is_feasible, double reginit);
CGX_API int cgx_solver_download(cgx_solver_t s, unsigned char* converged);

/* step-wise API (what unittest/bindings/test_solvers.py drives on the reference) */
/* SolverAbstract::setCandidate (solver-base.cpp:42-67); also resets iter_ = 0 and was_feasible_ = false */
CGX_API int cgx_solver_set_candidate(cgx_solver_t s, const double* xs, const double* us, int is_feasible);
/* SolverDDP::computeDirection(recalc) = calcDiff() + backwardPass() (ddp.cpp:120-125,157-253).
 * status[b] (may be NULL): 0 ok, 1 "backward_error" (the exception the reference throws, ddp.cpp:246-251,302-304) */
CGX_API int cgx_solver_compute_direction(cgx_solver_t s, int recalc, int* status);
/* SolverDDP::tryStep(alpha) (ddp.cpp:127-130) -> dV[b] = cost - cost_try after forwardPass(alpha)
 * (ddp.cpp:255-296 / fddp.cpp:149-225); status[b]: 0 ok, 1 "forward_error" */
CGX_API int cgx_solver_try_step(cgx_solver_t s, double alpha, double* dV, int* status);
/* expectedImprovement (ddp.cpp:144-155 / fddp.cpp:107-147, FDDP includes updateExpectedImprovement): d = [B][2] */
CGX_API int cgx_solver_expected_improvement(cgx_solver_t s, double* d);
/* stoppingCriteria (ddp.cpp:132-142): stop = [B] */
CGX_API int cgx_solver_stopping_criteria(cgx_solver_t s, double* stop);

/* accessors. name in {"xs","us","xs_try","us_try","fs","K","k","Vx","Vxx","Qxx","Qxu","Quu","Qx","Qu"}
 * (solver-base.hpp:162-167, ddp.hpp:191-236) and, for the box solvers, "Quu_inv" (box-ddp.hpp get_Quu_inv).
 * t == T is valid for xs, xs_try, fs, Vx, Vxx.
 * cgx_solver_get: one node of one problem, returns the number of doubles written.
 * cgx_solver_get_all: all nodes of all problems, out = [B][nt][size]. */
CGX_API int cgx_solver_get(cgx_solver_t s, const char* name, int b, int t, double* out);
CGX_API int cgx_solver_get_all(cgx_solver_t s, const char* name, double* out);
/* set_xs / set_us (solver-base.cpp:143-163) */
CGX_API int cgx_solver_set_xs(cgx_solver_t s, const double* xs);
CGX_API int cgx_solver_set_us(cgx_solver_t s, const double* us);

/* per-problem scalars, out = [B][CGX_NSCALARS] (solver-base.hpp:172-227):
 *  0 cost 1 stop 2 d[0] 3 d[1] 4 xreg 5 ureg 6 steplength 7 dV 8 dVexp 9 iter 10 is_feasible 11 converged
 *  12 n_records 13 cost_try 14 was_feasible 15 backward sweeps executed 16 forward sweeps executed
 *  (the sweep counts give the algorithmic HBM bytes of a solve) */
#define CGX_NSCALARS 17
CGX_API int cgx_solver_get_scalars(cgx_solver_t s, double* out);
/* the same for ONE problem: out = [CGX_NSCALARS] (the per-problem getters get_cost(), get_stop(), ... of the facade) */
CGX_API int cgx_solver_get_scalars_row(cgx_solver_t s, int b, double* out);

/* one row per iteration, out = [n][CGX_NTRACE]:
 *  0 iter 1 cost 2 stop 3 d0 4 d1 5 xreg 6 ureg 7 steplength 8 is_feasible 9 dV 10 dVexp
 *  11 index of the accepted step length (-1 none) 12 backward-pass retries
 * Returns the number of records of problem b (call with out == NULL to query). */
#define CGX_NTRACE 13
CGX_API int cgx_solver_get_trace(cgx_solver_t s, int b, double* out, int max_records);

/* CallbackAbstract (solver-base.hpp:284-298): fn is invoked on the calling host thread, once per recorded
 * iteration of every problem, after cgx_solver_solve finished the device loop (the loop itself never
 * returns to the host between iterations). Requires cgx_solver_enable_trace. */
typedef struct cgx_iter_record {
  int problem;
  int iter;
  double cost, stop, d0, d1, xreg, ureg, steplength, dV, dVexp;
  int is_feasible;
  int alpha_index;
  int retries;
} cgx_iter_record;
typedef void (*cgx_callback_fn)(void* user, const cgx_iter_record* rec);
CGX_API int cgx_solver_set_callback(cgx_solver_t s, cgx_callback_fn fn, void* user);

/* device timing of the last solve_resident / compute_direction / try_step call, CUDA events on the
 * solver's own stream: ms[0] = whole call, ms[1] = main solve/backward kernel, ms[2] = forward kernel;
 * launches = kernels launched by that call */
CGX_API int cgx_solver_last_timing(cgx_solver_t s, double* ms, int* launches);

/* FP64 DFMA throughput probe used as the roofline denominator for the Riccati kernels (SURVEY 8d:
 * the FP64 peak is not in MEASURED_PEAKS.json). Returns achieved TFLOP/s of a register-resident DFMA loop. */
CGX_API int cgx_probe_fp64_tflops(double* tflops);

/* ----------------------------------------------------------------------------------------------
 * cgx_unicycle_calc_noalloc — NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * ShootingProblem::calc for B unicycle problems (what cgx_problem_calc returns for CGX_MODEL_UNICYCLE: costs[b] = total
 * cost of xs[b], us[b]) without creating a handle and without any cudaMalloc / cudaFree / stream / event: the staging of
 * xs / us and the per-node costs between the node kernel and the reduction live in ONE static device array of
 * CGX_NOALLOC_SCRATCH_DOUBLES doubles (8 MiB per GPU that loads the library, held until exit).
 *   params  = [dt, w_x, w_u] of the running model, tparams = the terminal model's (NULL: the running model's);
 *   xs = [B][T+1][3], us = [B][T][2], costs = [B]; host or device pointers (copied with cudaMemcpyDefault);
 *   runs on the current device's legacy default stream and returns after the result has arrived.
 * Limit: B * (6 T + 5) <= CGX_NOALLOC_SCRATCH_DOUBLES (e.g. B <= 1733 at T = 100), else CGX_ERR_INVALID — nothing is
 * allocated to make room. Every call shares the array (it keeps the last call's contents; a call writes every word
 * before reading it, nothing needs clearing): calls that overlap on the GPU race on it — serialise them. */
#define CGX_NOALLOC_SCRATCH_DOUBLES 1048576
CGX_API int cgx_unicycle_calc_noalloc(int B, int T, const double* params, const double* tparams, const double* xs,
                                      const double* us, double* costs);

#ifdef __cplusplus
}
#endif
#endif /* CROCODDYL_B200_H_ */
