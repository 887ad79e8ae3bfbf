// crocoddyl_b200 — host-side backend interface behind the C ABI (include/crocoddyl_b200.h).
// Two backends implement it: thread-per-problem (tpp_backend.cu; unicycle) and CTA-per-problem
// (cta_backend.cu; LQR and the revolute-chain arm). All buffers named d_* are device pointers in the
// HOST ("user") layout of the C ABI: [B][nt][size], matrices column-major.
#pragma once
#include <cuda_runtime.h>
#include <stdexcept>
#include <string>
#include <vector>
#include "common.cuh"

namespace cgx {

struct CudaError : std::runtime_error {
  explicit CudaError(const std::string& w) : std::runtime_error(w) {}
};
struct InvalidArg : std::runtime_error {
  explicit InvalidArg(const std::string& w) : std::runtime_error(w) {}
};
struct Unsupported : std::runtime_error {
  explicit Unsupported(const std::string& w) : std::runtime_error(w) {}
};
struct NotFound : std::runtime_error {
  explicit NotFound(const std::string& w) : std::runtime_error(w) {}
};

#define CGX_CUDA(expr)                                                                                     \
  do {                                                                                                     \
    cudaError_t e_ = (expr);                                                                               \
    if (e_ != cudaSuccess)                                                                                 \
      throw cgx::CudaError(std::string(#expr) + ": " + cudaGetErrorString(e_) + " (" + __FILE__ + ":" +    \
                           std::to_string(__LINE__) + ")");                                               \
  } while (0)

template <class T>
inline T* dalloc(size_t n) {
  T* p = nullptr;
  if (n == 0) n = 1;
  CGX_CUDA(cudaMalloc((void**)&p, n * sizeof(T)));
  return p;
}

// The batch of ShootingProblems (device-resident description)
struct ProblemDesc {
  int kind, B, T, nx, nu;
  long nparams;
  int per_problem_params;
  std::vector<double> h_params;  // host copy (B blocks or 1)
  std::vector<double> h_x0;      // [B][nx]
  double* d_params = nullptr;    // device copy, same layout
  double* d_x0 = nullptr;        // [B][nx]
  // terminal model (shooting.hpp:48-63): same layout / count / per-problem-ness as the running block; by default a copy
  // of it (arm: with dt = 0, benchmark/factory/arm.hpp:89-90). Always allocated: kernels read it unconditionally.
  std::vector<double> h_tparams;
  double* d_tparams = nullptr;
  // control limits of the running model (action-base.hxx:117-140): [nb][nu] each, nb = B if bounds_per_problem else 1;
  // d_haslim[nb] = has_control_limits. Always allocated (defaults -inf / +inf / 0).
  int bounds_per_problem = 0;
  std::vector<double> h_ulb, h_uub;
  std::vector<int> h_haslim;
  double* d_ulb = nullptr;
  double* d_uub = nullptr;
  int* d_haslim = nullptr;
};

// field description for accessors
struct FieldInfo {
  int size;  // doubles per node
  int nt;    // number of nodes (T or T+1)
};

struct Backend {
  const ProblemDesc* p = nullptr;
  int solver_kind = 1;
  SolverOpts o;
  int full = 1;
  int trace_on = 0, trace_cap = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  double last_ms[3] = {0, 0, 0};
  int last_launches = 0;

  virtual ~Backend() {}
  // (re)allocate working storage after a change of storage policy / trace settings
  virtual void configure() = 0;
  virtual void x0_changed() = 0;
  // candidate from device staging buffers in user layout (null = zeros), solver-base.cpp:42-67
  virtual void set_candidate(const double* d_xs, const double* d_us, int is_feasible) = 0;
  virtual void solve(int maxiter, int is_feasible, double reginit) = 0;
  // Optional: the same solve with the candidate read from HOST buffers and xs/us (may be null) written back to HOST
  // buffers, in slices of the batch on separate streams so that the PCIe copies overlap the solve (thread-per-problem
  // backend). d_xs/d_us = device staging buffers in user layout. false = not available, nothing was done.
  virtual bool solve_pipelined(const double*, const double*, double*, double*, int, int, double, double*, double*) {
    return false;
  }
  virtual void compute_direction(int recalc) = 0;
  virtual void try_step(double alpha) = 0;
  // accessors: export problems [b0, b0+nb) of a field to d_out ([nb][nt][size], user layout)
  virtual FieldInfo field_info(const std::string& name) = 0;
  virtual void export_field(const std::string& name, int b0, int nb, double* d_out) = 0;
  virtual void import_field(const std::string& name, const double* d_in) = 0;  // xs / us only
  // scalars: d_out = [B][17] as documented in the C header; statuses: [B] ints
  virtual void export_scalars(double* d_out) = 0;
  virtual void export_status(int which /*0 backward, 1 forward*/, int* d_out) = 0;
  virtual void set_reg(double xreg, double ureg, bool set_x, bool set_u) = 0;
  virtual void export_trace(int b, double* d_out /*[cap][NTRACE]*/) = 0;
  // records [0, nrec) of every problem in one go: d_out = [B][nrec][NTRACE] (nrec <= trace_cap)
  virtual void export_trace_all(int nrec, double* d_out) = 0;
  // problem-level API (ShootingProblem::calc / calcDiff / rollout / quasiStatic)
  virtual void problem_calc(const double* d_xs, const double* d_us, int diff, double* d_costs) = 0;
  virtual void problem_rollout(const double* d_us, double* d_xs) = 0;
  virtual void problem_quasi_static(const double* d_x, double* d_u) = 0;
  virtual FieldInfo data_info(const std::string& name) = 0;
  virtual void export_data(const std::string& name, int b, int t, double* d_out) = 0;
};

Backend* make_tpp_backend(const ProblemDesc* p, int solver_kind);
Backend* make_cta_backend(const ProblemDesc* p, int solver_kind);

}  // namespace cgx
