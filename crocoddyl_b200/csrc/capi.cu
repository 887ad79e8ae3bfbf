// crocoddyl_b200 — extern "C" layer (include/crocoddyl_b200.h) over the CUDA backends.
// No CPU fallback: every compute entry point needs a CUDA device and fails with CGX_ERR_CUDA otherwise.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <string>
#include "../../include/crocoddyl_b200.h"
#include "backend.hpp"

using namespace cgx;

struct cgx_problem_s {
  ProblemDesc d;
  Backend* pb = nullptr;  // backend instance serving the problem-level API (calc/calcDiff/rollout)
  std::vector<Backend*> live;  // backends of the solvers created on this problem (told about set_x0)
  double* d_tmp_xs = nullptr;
  double* d_tmp_us = nullptr;
  double* d_tmp_cost = nullptr;
  double* d_tmp_data = nullptr;  // one ActionData field of one node (largest: ndx x ndx)
};

// page-locked host block for the per-problem scalar record: the D2H copy of [B][17] doubles runs at PCIe rate instead
// of through the driver's pageable staging (8.9 MB at B = 65,536, fetched after every solve)
struct PinnedDoubles {
  double* p = nullptr;
  size_t n = 0, cap = 0;
  ~PinnedDoubles() {
    if (p) cudaFreeHost(p);
  }
  void resize(size_t m) {
    if (m > cap) {
      if (p) cudaFreeHost(p);
      p = nullptr;
      cap = 0;
      if (cudaMallocHost((void**)&p, m * sizeof(double)) != cudaSuccess) {
        cudaGetLastError();
        throw CudaError("cudaMallocHost failed for the scalar record");
      }
      cap = m;
    }
    n = m;
  }
  double* data() { return p; }
  size_t size() const { return n; }
  double& operator[](size_t i) { return p[i]; }
};

struct cgx_solver_s {
  cgx_problem_s* p = nullptr;
  Backend* be = nullptr;
  double* d_init_xs = nullptr;  // staging, user layout
  double* d_init_us = nullptr;
  bool has_xs = false, has_us = false;
  double* d_scal = nullptr;  // [B][17]
  PinnedDoubles h_scal;
  PinnedDoubles h_trace;  // [B][nrec][NTRACE] for the callback replay
  int* d_status = nullptr;
  double* d_field = nullptr;  // scratch for accessors
  size_t field_cap = 0;
  cgx_callback_fn cb = nullptr;
  void* cb_user = nullptr;
};

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}

#define CGX_TRY try {
#define CGX_CATCH                                                   \
  }                                                                 \
  catch (const CudaError& e) { return fail(CGX_ERR_CUDA, e.what()); } \
  catch (const InvalidArg& e) { return fail(CGX_ERR_INVALID, e.what()); } \
  catch (const Unsupported& e) { return fail(CGX_ERR_UNSUPPORTED, e.what()); } \
  catch (const NotFound& e) { return fail(CGX_ERR_NOTFOUND, e.what()); } \
  catch (const std::exception& e) { return fail(CGX_ERR_INVALID, e.what()); }

void require_device() {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    cudaGetLastError();
    throw CudaError("no CUDA device available: crocoddyl_b200 has no CPU fallback");
  }
}

size_t lqr_param_count(int n, int m) { return 1 + 2 * (size_t)n * n + 2 * (size_t)n * m + (size_t)m * m + 2 * n + m; }

Backend* make_backend(const ProblemDesc* d, int solver_kind) {
  if (d->kind == CGX_MODEL_UNICYCLE) return make_tpp_backend(d, solver_kind);
  return make_cta_backend(d, solver_kind);
}

void default_opts(SolverOpts& o) {
  // solver-base.cpp:24-25, ddp.cpp:17-36, fddp.cpp:15
  o.th_acceptstep = 0.1;
  o.th_stop = 1e-9;
  o.th_grad = 1e-12;
  o.th_stepdec = 0.5;
  o.th_stepinc = 0.01;
  o.th_acceptnegstep = 2.;
  o.regfactor = 10.;
  o.regmin = 1e-9;
  o.regmax = 1e9;
  o.nalphas = 10;
  for (int n = 0; n < 10; ++n) o.alphas[n] = 1. / std::pow(2., (double)n);
  if (o.th_stepinc < o.alphas[o.nalphas - 1]) o.th_stepinc = o.alphas[o.nalphas - 1];
}

void finish_timing(Backend* be) {
  CGX_CUDA(cudaEventRecord(be->ev[3], be->stream));
  CGX_CUDA(cudaStreamSynchronize(be->stream));
  float ms = 0;
  CGX_CUDA(cudaEventElapsedTime(&ms, be->ev[0], be->ev[3]));
  be->last_ms[0] = ms;
  if (cudaEventElapsedTime(&ms, be->ev[1], be->ev[2]) == cudaSuccess)
    be->last_ms[1] = ms;
  else
    cudaGetLastError();
}
void start_timing(Backend* be) {
  be->last_launches = 0;
  be->last_ms[0] = be->last_ms[1] = be->last_ms[2] = 0;
  CGX_CUDA(cudaEventRecord(be->ev[0], be->stream));
}

double* scratch(cgx_solver_s* s, size_t n) {
  if (n > s->field_cap) {
    if (s->d_field) cudaFree(s->d_field);
    s->d_field = dalloc<double>(n);
    s->field_cap = n;
  }
  return s->d_field;
}

void fetch_scalars(cgx_solver_s* s) {
  const int B = s->p->d.B;
  s->be->export_scalars(s->d_scal);
  s->h_scal.resize((size_t)B * 17);
  CGX_CUDA(cudaMemcpyAsync(s->h_scal.data(), s->d_scal, s->h_scal.size() * sizeof(double), cudaMemcpyDeviceToHost,
                           s->be->stream));
  CGX_CUDA(cudaStreamSynchronize(s->be->stream));
}

}  // namespace

// for the other translation units of the library (noalloc.cu): store the message cgx_last_error() returns
namespace cgx {
int set_last_error(int code, const std::string& msg) { return fail(code, msg); }
}  // namespace cgx

extern "C" {

const char* cgx_last_error(void) { return g_err.c_str(); }
const char* cgx_version(void) { return "crocoddyl_b200 0.1 (sm_100a)"; }
int cgx_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}
int cgx_set_device(int device) {
  CGX_TRY
  require_device();
  CGX_CUDA(cudaSetDevice(device));
  return CGX_OK;
  CGX_CATCH
}

long cgx_arm_param_count(int nv) { return 36 + 28L * nv; }

int cgx_problem_create(cgx_problem_t* out, int kind, int B, int T, int nx, int nu, const double* params,
                       long nparams, int per_problem_params, const double* x0s) {
  CGX_TRY
  if (!out) throw InvalidArg("Invalid argument: out is NULL");
  *out = nullptr;
  if (B <= 0 || T <= 0 || nx <= 0 || nu < 0) throw InvalidArg("Invalid argument: B, T, nx must be positive");
  if (!params || !x0s) throw InvalidArg("Invalid argument: params / x0s is NULL");
  if (kind == CGX_MODEL_LQR) {
    if ((size_t)nparams != lqr_param_count(nx, nu)) throw InvalidArg("Invalid argument: LQR params have wrong size");
  } else if (kind == CGX_MODEL_UNICYCLE) {
    if (nparams != 3 || nx != 3 || nu != 2) throw InvalidArg("Invalid argument: unicycle needs nx=3, nu=2, 3 params");
  } else if (kind == CGX_MODEL_ARM) {
    if (nx % 2 != 0 || nu != nx / 2 || nparams != cgx_arm_param_count(nu))
      throw InvalidArg("Invalid argument: arm needs nx = 2 nv, nu = nv and cgx_arm_param_count(nv) params");
  } else {
    throw InvalidArg("Invalid argument: unknown model kind");
  }
  require_device();
  cgx_problem_s* p = new cgx_problem_s();
  p->d.kind = kind;
  p->d.B = B;
  p->d.T = T;
  p->d.nx = nx;
  p->d.nu = nu;
  p->d.nparams = nparams;
  p->d.per_problem_params = per_problem_params != 0;
  const size_t np = (size_t)nparams * (per_problem_params ? B : 1);
  p->d.h_params.assign(params, params + np);
  p->d.h_x0.assign(x0s, x0s + (size_t)B * nx);
  try {
    p->d.d_params = dalloc<double>(np);
    p->d.d_x0 = dalloc<double>((size_t)B * nx);
    CGX_CUDA(cudaMemcpy(p->d.d_params, params, np * sizeof(double), cudaMemcpyHostToDevice));
    CGX_CUDA(cudaMemcpy(p->d.d_x0, x0s, (size_t)B * nx * sizeof(double), cudaMemcpyHostToDevice));
    // terminal block: the running one (arm: with dt = 0, benchmark/factory/arm.hpp:89-90) until
    // cgx_problem_set_terminal_params replaces it
    p->d.h_tparams = p->d.h_params;
    if (kind == CGX_MODEL_ARM)
      for (size_t b = 0; b < (per_problem_params ? (size_t)B : 1); ++b) p->d.h_tparams[b * nparams] = 0.;
    p->d.d_tparams = dalloc<double>(np);
    CGX_CUDA(cudaMemcpy(p->d.d_tparams, p->d.h_tparams.data(), np * sizeof(double), cudaMemcpyHostToDevice));
    // control limits: unbounded (action-base.hxx:19-22)
    const size_t nb = (size_t)B * (nu ? nu : 1);
    p->d.h_ulb.assign(nb, -std::numeric_limits<double>::infinity());
    p->d.h_uub.assign(nb, std::numeric_limits<double>::infinity());
    p->d.h_haslim.assign(B, 0);
    p->d.d_ulb = dalloc<double>(nb);
    p->d.d_uub = dalloc<double>(nb);
    p->d.d_haslim = dalloc<int>(B);
    CGX_CUDA(cudaMemcpy(p->d.d_ulb, p->d.h_ulb.data(), nb * sizeof(double), cudaMemcpyHostToDevice));
    CGX_CUDA(cudaMemcpy(p->d.d_uub, p->d.h_uub.data(), nb * sizeof(double), cudaMemcpyHostToDevice));
    CGX_CUDA(cudaMemset(p->d.d_haslim, 0, B * sizeof(int)));
  } catch (...) {
    cgx_problem_destroy(p);
    throw;
  }
  *out = p;
  return CGX_OK;
  CGX_CATCH
}

int cgx_problem_destroy(cgx_problem_t p) {
  if (!p) return CGX_OK;
  delete p->pb;
  if (p->d.d_params) cudaFree(p->d.d_params);
  if (p->d.d_x0) cudaFree(p->d.d_x0);
  if (p->d.d_tparams) cudaFree(p->d.d_tparams);
  if (p->d.d_ulb) cudaFree(p->d.d_ulb);
  if (p->d.d_uub) cudaFree(p->d.d_uub);
  if (p->d.d_haslim) cudaFree(p->d.d_haslim);
  if (p->d_tmp_xs) cudaFree(p->d_tmp_xs);
  if (p->d_tmp_us) cudaFree(p->d_tmp_us);
  if (p->d_tmp_cost) cudaFree(p->d_tmp_cost);
  if (p->d_tmp_data) cudaFree(p->d_tmp_data);
  delete p;
  return CGX_OK;
}

int cgx_problem_set_terminal_params(cgx_problem_t p, const double* tparams, long nparams) {
  CGX_TRY
  if (!p || !tparams) throw InvalidArg("Invalid argument: NULL");
  if (nparams != p->d.nparams)
    throw InvalidArg("Invalid argument: terminal params have wrong size (it should be " + std::to_string(p->d.nparams) + ")");
  const size_t np = (size_t)nparams * (p->d.per_problem_params ? p->d.B : 1);
  p->d.h_tparams.assign(tparams, tparams + np);
  CGX_CUDA(cudaDeviceSynchronize());  // no kernel of a live solver may be reading the old block
  CGX_CUDA(cudaMemcpy(p->d.d_tparams, tparams, np * sizeof(double), cudaMemcpyHostToDevice));
  return CGX_OK;
  CGX_CATCH
}
int cgx_problem_get_terminal_params(cgx_problem_t p, double* tparams) {
  if (!p || !tparams) return fail(CGX_ERR_INVALID, "Invalid argument: NULL");
  std::memcpy(tparams, p->d.h_tparams.data(), p->d.h_tparams.size() * sizeof(double));
  return CGX_OK;
}

int cgx_problem_set_u_bounds(cgx_problem_t p, const double* lb, const double* ub, int per_problem) {
  CGX_TRY
  if (!p || !lb || !ub) throw InvalidArg("Invalid argument: NULL");
  const int B = p->d.B, nu = p->d.nu;
  if (nu == 0) throw InvalidArg("Invalid argument: lower bound has wrong dimension (it should be 0)");
  for (int b = 0; b < B; ++b) {
    const double* l = lb + (per_problem ? (size_t)b * nu : 0);
    const double* u = ub + (per_problem ? (size_t)b * nu : 0);
    bool fl = false, fu = false;
    for (int i = 0; i < nu; ++i) {
      if (std::isnan(l[i]) || std::isnan(u[i])) throw InvalidArg("Invalid argument: NaN in the control bounds");
      p->d.h_ulb[(size_t)b * nu + i] = l[i];
      p->d.h_uub[(size_t)b * nu + i] = u[i];
      fl = fl || std::isfinite(l[i]);
      fu = fu || std::isfinite(u[i]);
    }
    p->d.h_haslim[b] = (fl && fu) ? 1 : 0;  // action-base.hxx:142-144
  }
  p->d.bounds_per_problem = per_problem != 0;
  CGX_CUDA(cudaDeviceSynchronize());
  CGX_CUDA(cudaMemcpy(p->d.d_ulb, p->d.h_ulb.data(), (size_t)B * nu * sizeof(double), cudaMemcpyHostToDevice));
  CGX_CUDA(cudaMemcpy(p->d.d_uub, p->d.h_uub.data(), (size_t)B * nu * sizeof(double), cudaMemcpyHostToDevice));
  CGX_CUDA(cudaMemcpy(p->d.d_haslim, p->d.h_haslim.data(), (size_t)B * sizeof(int), cudaMemcpyHostToDevice));
  return CGX_OK;
  CGX_CATCH
}
int cgx_problem_get_u_bounds(cgx_problem_t p, double* lb, double* ub, int* per_problem, int* has_control_limits) {
  if (!p) return fail(CGX_ERR_INVALID, "Invalid argument: NULL");
  const size_t cnt = (size_t)(p->d.bounds_per_problem ? p->d.B : 1) * p->d.nu;
  if (lb) std::memcpy(lb, p->d.h_ulb.data(), cnt * sizeof(double));
  if (ub) std::memcpy(ub, p->d.h_uub.data(), cnt * sizeof(double));
  if (per_problem) *per_problem = p->d.bounds_per_problem;
  if (has_control_limits)
    for (int b = 0; b < (p->d.bounds_per_problem ? p->d.B : 1); ++b) has_control_limits[b] = p->d.h_haslim[b];
  return CGX_OK;
}

int cgx_problem_dims(cgx_problem_t p, int* B, int* T, int* nx, int* ndx, int* nu) {
  if (!p) return fail(CGX_ERR_INVALID, "Invalid argument: problem is NULL");
  if (B) *B = p->d.B;
  if (T) *T = p->d.T;
  if (nx) *nx = p->d.nx;
  if (ndx) *ndx = p->d.nx;
  if (nu) *nu = p->d.nu;
  return CGX_OK;
}

int cgx_problem_get_x0(cgx_problem_t p, double* x0s) {
  if (!p || !x0s) return fail(CGX_ERR_INVALID, "Invalid argument: NULL");
  std::memcpy(x0s, p->d.h_x0.data(), p->d.h_x0.size() * sizeof(double));
  return CGX_OK;
}

static void problem_backend(cgx_problem_s* p) {
  if (!p->pb) {
    p->pb = make_backend(&p->d, CGX_SOLVER_FDDP);
    const size_t B = p->d.B, T = p->d.T;
    p->d_tmp_xs = dalloc<double>(B * (T + 1) * p->d.nx);
    p->d_tmp_us = dalloc<double>(B * T * (p->d.nu ? p->d.nu : 1));
    p->d_tmp_cost = dalloc<double>(B);
    const size_t n = p->d.nx, m = p->d.nu ? p->d.nu : 1;
    p->d_tmp_data = dalloc<double>(n * n + n * m + m * m);
  }
}

int cgx_problem_set_x0(cgx_problem_t p, const double* x0s) {
  CGX_TRY
  if (!p || !x0s) throw InvalidArg("Invalid argument: NULL");
  p->d.h_x0.assign(x0s, x0s + (size_t)p->d.B * p->d.nx);
  CGX_CUDA(cudaMemcpy(p->d.d_x0, x0s, p->d.h_x0.size() * sizeof(double), cudaMemcpyHostToDevice));
  if (p->pb) p->pb->x0_changed();
  // every live solver keeps (or aliases) its own view of x0: shooting.hxx:393-400 changes what the NEXT solve() sees
  for (Backend* be : p->live) be->x0_changed();
  return CGX_OK;
  CGX_CATCH
}

static int problem_calc_impl(cgx_problem_t p, const double* xs, const double* us, double* costs, int diff) {
  CGX_TRY
  if (!p || !xs || !us) throw InvalidArg("Invalid argument: xs / us is NULL");
  problem_backend(p);
  Backend* be = p->pb;
  const size_t B = p->d.B, T = p->d.T;
  CGX_CUDA(cudaMemcpyAsync(p->d_tmp_xs, xs, B * (T + 1) * p->d.nx * sizeof(double), cudaMemcpyHostToDevice, be->stream));
  CGX_CUDA(cudaMemcpyAsync(p->d_tmp_us, us, B * T * p->d.nu * sizeof(double), cudaMemcpyHostToDevice, be->stream));
  be->problem_calc(p->d_tmp_xs, p->d_tmp_us, diff, p->d_tmp_cost);
  if (costs) CGX_CUDA(cudaMemcpyAsync(costs, p->d_tmp_cost, B * sizeof(double), cudaMemcpyDeviceToHost, be->stream));
  CGX_CUDA(cudaStreamSynchronize(be->stream));
  CGX_CUDA(cudaGetLastError());
  return CGX_OK;
  CGX_CATCH
}
int cgx_problem_calc(cgx_problem_t p, const double* xs, const double* us, double* costs) {
  return problem_calc_impl(p, xs, us, costs, 0);
}
int cgx_problem_calc_diff(cgx_problem_t p, const double* xs, const double* us, double* costs) {
  return problem_calc_impl(p, xs, us, costs, 1);
}
int cgx_problem_rollout(cgx_problem_t p, const double* us, double* xs) {
  CGX_TRY
  if (!p || !xs || !us) throw InvalidArg("Invalid argument: xs / us is NULL");
  problem_backend(p);
  Backend* be = p->pb;
  const size_t B = p->d.B, T = p->d.T;
  CGX_CUDA(cudaMemcpyAsync(p->d_tmp_us, us, B * T * p->d.nu * sizeof(double), cudaMemcpyHostToDevice, be->stream));
  be->problem_rollout(p->d_tmp_us, p->d_tmp_xs);
  CGX_CUDA(cudaMemcpyAsync(xs, p->d_tmp_xs, B * (T + 1) * p->d.nx * sizeof(double), cudaMemcpyDeviceToHost, be->stream));
  CGX_CUDA(cudaStreamSynchronize(be->stream));
  CGX_CUDA(cudaGetLastError());
  return CGX_OK;
  CGX_CATCH
}
int cgx_problem_quasi_static(cgx_problem_t p, const double* xs_point, double* us_out) {
  CGX_TRY
  if (!p || !xs_point || !us_out) throw InvalidArg("Invalid argument: NULL");
  problem_backend(p);
  Backend* be = p->pb;
  const size_t B = p->d.B;
  CGX_CUDA(cudaMemcpyAsync(p->d_tmp_xs, xs_point, B * p->d.nx * sizeof(double), cudaMemcpyHostToDevice, be->stream));
  be->problem_quasi_static(p->d_tmp_xs, p->d_tmp_us);
  CGX_CUDA(cudaMemcpyAsync(us_out, p->d_tmp_us, B * p->d.nu * sizeof(double), cudaMemcpyDeviceToHost, be->stream));
  CGX_CUDA(cudaStreamSynchronize(be->stream));
  CGX_CUDA(cudaGetLastError());
  return CGX_OK;
  CGX_CATCH
}
int cgx_problem_get_data(cgx_problem_t p, const char* field, int b, int t, double* out) {
  CGX_TRY
  if (!p || !field || !out) throw InvalidArg("Invalid argument: NULL");
  if (b < 0 || b >= p->d.B || t < 0 || t > p->d.T) throw InvalidArg("Invalid argument: b / t out of range");
  problem_backend(p);
  Backend* be = p->pb;
  FieldInfo fi = be->data_info(field);
  be->export_data(field, b, t, p->d_tmp_data);
  CGX_CUDA(cudaMemcpyAsync(out, p->d_tmp_data, fi.size * sizeof(double), cudaMemcpyDeviceToHost, be->stream));
  CGX_CUDA(cudaStreamSynchronize(be->stream));
  return fi.size;
  CGX_CATCH
}

// ------------------------------------------------------------------------------------------------
int cgx_solver_create(cgx_solver_t* out, cgx_problem_t p, int solver_kind) {
  CGX_TRY
  if (!out || !p) throw InvalidArg("Invalid argument: NULL");
  *out = nullptr;
  if (solver_kind < CGX_SOLVER_DDP || solver_kind > CGX_SOLVER_BOXFDDP) throw InvalidArg("Invalid argument: solver kind");
  require_device();
  cgx_solver_s* s = new cgx_solver_s();
  s->p = p;
  try {
    s->be = make_backend(&p->d, solver_kind);
    default_opts(s->be->o);
    // box-ddp.cpp:24-29 / box-fddp.cpp:24-29: the gradient of the Lagrangian is not available, looser default tolerance
    if (solver_kind == CGX_SOLVER_BOXDDP || solver_kind == CGX_SOLVER_BOXFDDP) s->be->o.th_stop = 5e-5;
    s->be->configure();
    const size_t B = p->d.B, T = p->d.T;
    s->d_init_xs = dalloc<double>(B * (T + 1) * p->d.nx);
    s->d_init_us = dalloc<double>(B * T * (p->d.nu ? p->d.nu : 1));
    s->d_scal = dalloc<double>(B * 17);
    s->d_status = dalloc<int>(B);
  } catch (...) {
    cgx_solver_destroy(s);
    throw;
  }
  p->live.push_back(s->be);
  *out = s;
  return CGX_OK;
  CGX_CATCH
}

int cgx_solver_destroy(cgx_solver_t s) {
  if (!s) return CGX_OK;
  if (s->p && s->be)
    for (size_t i = 0; i < s->p->live.size(); ++i)
      if (s->p->live[i] == s->be) {
        s->p->live.erase(s->p->live.begin() + i);
        break;
      }
  delete s->be;
  if (s->d_init_xs) cudaFree(s->d_init_xs);
  if (s->d_init_us) cudaFree(s->d_init_us);
  if (s->d_scal) cudaFree(s->d_scal);
  if (s->d_status) cudaFree(s->d_status);
  if (s->d_field) cudaFree(s->d_field);
  delete s;
  return CGX_OK;
}

int cgx_solver_set_param(cgx_solver_t s, const char* name, double v) {
  CGX_TRY
  if (!s || !name) throw InvalidArg("Invalid argument: NULL");
  SolverOpts& o = s->be->o;
  const std::string n(name);
  // value checks as the reference setters (solver-base.cpp:165-173, ddp.cpp:420-486, fddp.cpp:229-235)
  if (n == "th_acceptstep") {
    if (!(0. < v && v <= 1.)) throw InvalidArg("Invalid argument: th_acceptstep value should between 0 and 1.");
    o.th_acceptstep = v;
  } else if (n == "th_stop") {
    if (!(v > 0.)) throw InvalidArg("Invalid argument: th_stop value has to higher than 0.");
    o.th_stop = v;
  } else if (n == "th_grad") {
    if (v < 0.) throw InvalidArg("Invalid argument: th_grad value has to be positive.");
    o.th_grad = v;
  } else if (n == "th_stepdec") {
    if (!(0. < v && v <= 1.)) throw InvalidArg("Invalid argument: th_stepdec value should between 0 and 1.");
    o.th_stepdec = v;
  } else if (n == "th_stepinc") {
    if (!(0. < v && v <= 1.)) throw InvalidArg("Invalid argument: th_stepinc value should between 0 and 1.");
    o.th_stepinc = v;
  } else if (n == "th_acceptnegstep") {
    if (s->be->solver_kind != CGX_SOLVER_FDDP && s->be->solver_kind != CGX_SOLVER_BOXFDDP)
      throw InvalidArg("Invalid argument: th_acceptnegstep is an FDDP parameter");
    if (v < 0.) throw InvalidArg("Invalid argument: th_acceptnegstep value has to be positive.");
    o.th_acceptnegstep = v;
  } else if (n == "regfactor") {
    if (v <= 1.) throw InvalidArg("Invalid argument: regfactor value is higher than 1.");
    o.regfactor = v;
  } else if (n == "regmin") {
    if (v < 0.) throw InvalidArg("Invalid argument: regmin value has to be positive.");
    o.regmin = v;
  } else if (n == "regmax") {
    if (v < 0.) throw InvalidArg("Invalid argument: regmax value has to be positive.");
    o.regmax = v;
  } else if (n == "xreg" || n == "ureg") {
    if (v < 0.) throw InvalidArg("Invalid argument: " + n + " value has to be positive.");
    s->be->set_reg(v, v, n == "xreg", n == "ureg");
    CGX_CUDA(cudaStreamSynchronize(s->be->stream));
  } else {
    throw NotFound("unknown parameter " + n);
  }
  return CGX_OK;
  CGX_CATCH
}

int cgx_solver_get_param(cgx_solver_t s, const char* name, double* v) {
  CGX_TRY
  if (!s || !name || !v) throw InvalidArg("Invalid argument: NULL");
  const SolverOpts& o = s->be->o;
  const std::string n(name);
  if (n == "th_acceptstep") *v = o.th_acceptstep;
  else if (n == "th_stop") *v = o.th_stop;
  else if (n == "th_grad") *v = o.th_grad;
  else if (n == "th_stepdec") *v = o.th_stepdec;
  else if (n == "th_stepinc") *v = o.th_stepinc;
  else if (n == "th_acceptnegstep") *v = o.th_acceptnegstep;
  else if (n == "regfactor") *v = o.regfactor;
  else if (n == "regmin") *v = o.regmin;
  else if (n == "regmax") *v = o.regmax;
  else throw NotFound("unknown parameter " + n);
  return CGX_OK;
  CGX_CATCH
}

int cgx_solver_set_alphas(cgx_solver_t s, const double* alphas, int n) {
  CGX_TRY
  if (!s || !alphas) throw InvalidArg("Invalid argument: NULL");
  if (n < 1 || n > 16) throw InvalidArg("Invalid argument: between 1 and 16 step lengths are supported");
  double prev = alphas[0];
  if (prev != 1.) std::fprintf(stderr, "Warning: alpha[0] should be 1\n");  // ddp.cpp:449-451
  for (int i = 1; i < n; ++i) {
    if (alphas[i] <= 0.) throw InvalidArg("Invalid argument: alpha values has to be positive.");
    if (alphas[i] >= prev) throw InvalidArg("Invalid argument: alpha values are monotonously decreasing.");
    prev = alphas[i];
  }
  for (int i = 0; i < n; ++i) s->be->o.alphas[i] = alphas[i];
  s->be->o.nalphas = n;
  return CGX_OK;
  CGX_CATCH
}
int cgx_solver_get_alphas(cgx_solver_t s, double* alphas, int* n) {
  if (!s || !alphas || !n) return fail(CGX_ERR_INVALID, "Invalid argument: NULL");
  for (int i = 0; i < s->be->o.nalphas; ++i) alphas[i] = s->be->o.alphas[i];
  *n = s->be->o.nalphas;
  return CGX_OK;
}

int cgx_solver_set_storage(cgx_solver_t s, int full) {
  CGX_TRY
  if (!s) throw InvalidArg("Invalid argument: NULL");
  s->be->full = full != 0;
  s->be->configure();
  return CGX_OK;
  CGX_CATCH
}
int cgx_solver_enable_trace(cgx_solver_t s, int on, int max_records) {
  CGX_TRY
  if (!s) throw InvalidArg("Invalid argument: NULL");
  if (on && max_records <= 0) throw InvalidArg("Invalid argument: max_records must be positive");
  s->be->trace_on = on != 0;
  s->be->trace_cap = on ? max_records : 0;
  s->be->configure();
  return CGX_OK;
  CGX_CATCH
}

int cgx_solver_upload(cgx_solver_t s, const double* init_xs, const double* init_us) {
  CGX_TRY
  if (!s) throw InvalidArg("Invalid argument: NULL");
  const size_t B = s->p->d.B, T = s->p->d.T;
  s->has_xs = init_xs != nullptr;
  s->has_us = init_us != nullptr;
  if (init_xs)
    CGX_CUDA(cudaMemcpyAsync(s->d_init_xs, init_xs, B * (T + 1) * s->p->d.nx * sizeof(double), cudaMemcpyHostToDevice,
                             s->be->stream));
  if (init_us)
    CGX_CUDA(cudaMemcpyAsync(s->d_init_us, init_us, B * T * s->p->d.nu * sizeof(double), cudaMemcpyHostToDevice,
                             s->be->stream));
  CGX_CUDA(cudaStreamSynchronize(s->be->stream));
  return CGX_OK;
  CGX_CATCH
}

int cgx_solver_solve_resident(cgx_solver_t s, int maxiter, int is_feasible, double reginit) {
  CGX_TRY
  if (!s) throw InvalidArg("Invalid argument: NULL");
  if (maxiter < 0) throw InvalidArg("Invalid argument: maxiter");
  Backend* be = s->be;
  start_timing(be);
  be->set_candidate(s->has_xs ? s->d_init_xs : nullptr, s->has_us ? s->d_init_us : nullptr, is_feasible != 0);
  be->solve(maxiter, is_feasible != 0, reginit);
  finish_timing(be);
  CGX_CUDA(cudaGetLastError());
  return CGX_OK;
  CGX_CATCH
}

int cgx_solver_download(cgx_solver_t s, unsigned char* converged) {
  CGX_TRY
  if (!s) throw InvalidArg("Invalid argument: NULL");
  fetch_scalars(s);
  const int B = s->p->d.B;
  if (converged)
    for (int b = 0; b < B; ++b) converged[b] = s->h_scal[(size_t)b * 17 + 11] != 0.;
  return CGX_OK;
  CGX_CATCH
}

// 1 = done by the backend's sliced host pipeline, 0 = not available (nothing done), < 0 = error code
static int solve_pipelined(cgx_solver_t s, const double* init_xs, const double* init_us, int maxiter, int is_feasible,
                           double reginit, double* xs_out, double* us_out) {
  CGX_TRY
  if (!s) throw InvalidArg("Invalid argument: NULL");
  if (maxiter < 0) throw InvalidArg("Invalid argument: maxiter");
  if (!init_xs || !init_us) return 0;
  Backend* be = s->be;
  start_timing(be);
  if (!be->solve_pipelined(init_xs, init_us, s->d_init_xs, s->d_init_us, maxiter, is_feasible != 0, reginit, xs_out,
                           us_out))
    return 0;
  s->has_xs = s->has_us = true;
  finish_timing(be);
  CGX_CUDA(cudaGetLastError());
  return 1;
  CGX_CATCH
}

int cgx_solver_solve(cgx_solver_t s, const double* init_xs, const double* init_us, int maxiter, int is_feasible,
                     double reginit, unsigned char* converged) {
  return cgx_solver_solve_fetch(s, init_xs, init_us, maxiter, is_feasible, reginit, converged, nullptr, nullptr);
}

int cgx_solver_solve_fetch(cgx_solver_t s, const double* init_xs, const double* init_us, int maxiter, int is_feasible,
                           double reginit, unsigned char* converged, double* xs_out, double* us_out) {
  int rc = solve_pipelined(s, init_xs, init_us, maxiter, is_feasible, reginit, xs_out, us_out);
  if (rc < 0) return rc;
  if (rc == 0) {
    rc = cgx_solver_upload(s, init_xs, init_us);
    if (rc) return rc;
    rc = cgx_solver_solve_resident(s, maxiter, is_feasible, reginit);
    if (rc) return rc;
    if (xs_out && (rc = cgx_solver_get_all(s, "xs", xs_out))) return rc;
    if (us_out && (rc = cgx_solver_get_all(s, "us", us_out))) return rc;
  }
  rc = cgx_solver_download(s, converged);
  if (rc) return rc;
  if (s->cb) {
    CGX_TRY
    if (!s->be->trace_on) throw InvalidArg("callbacks need cgx_solver_enable_trace");
    // the whole trace block comes back with one kernel and one copy into page-locked memory, then is replayed here
    const int B = s->p->d.B;
    int nrec = 0;
    for (int b = 0; b < B; ++b) nrec = std::max(nrec, (int)s->h_scal[(size_t)b * 17 + 12]);
    nrec = std::min(nrec, s->be->trace_cap);
    if (nrec > 0) {
      const size_t cnt = (size_t)B * nrec * NTRACE;
      double* d = scratch(s, cnt);
      s->h_trace.resize(cnt);
      s->be->export_trace_all(nrec, d);
      CGX_CUDA(cudaMemcpyAsync(s->h_trace.data(), d, cnt * sizeof(double), cudaMemcpyDeviceToHost, s->be->stream));
      CGX_CUDA(cudaStreamSynchronize(s->be->stream));
      for (int b = 0; b < B; ++b) {
        const int n = std::min((int)s->h_scal[(size_t)b * 17 + 12], nrec);
        for (int i = 0; i < n; ++i) {
          const double* r = s->h_trace.data() + ((size_t)b * nrec + i) * NTRACE;
          cgx_iter_record ir;
          ir.problem = b;
          ir.iter = (int)r[0];
          ir.cost = r[1];
          ir.stop = r[2];
          ir.d0 = r[3];
          ir.d1 = r[4];
          ir.xreg = r[5];
          ir.ureg = r[6];
          ir.steplength = r[7];
          ir.is_feasible = (int)r[8];
          ir.dV = r[9];
          ir.dVexp = r[10];
          ir.alpha_index = (int)r[11];
          ir.retries = (int)r[12];
          s->cb(s->cb_user, &ir);
        }
      }
    }
    CGX_CATCH
  }
  return CGX_OK;
}

int cgx_solver_set_candidate(cgx_solver_t s, const double* xs, const double* us, int is_feasible) {
  int rc = cgx_solver_upload(s, xs, us);
  if (rc) return rc;
  CGX_TRY
  s->be->set_candidate(s->has_xs ? s->d_init_xs : nullptr, s->has_us ? s->d_init_us : nullptr, is_feasible != 0);
  CGX_CUDA(cudaStreamSynchronize(s->be->stream));
  CGX_CUDA(cudaGetLastError());
  return CGX_OK;
  CGX_CATCH
}

int cgx_solver_compute_direction(cgx_solver_t s, int recalc, int* status) {
  CGX_TRY
  if (!s) throw InvalidArg("Invalid argument: NULL");
  Backend* be = s->be;
  start_timing(be);
  be->compute_direction(recalc != 0);
  finish_timing(be);
  CGX_CUDA(cudaGetLastError());
  if (status) {
    be->export_status(0, s->d_status);
    CGX_CUDA(cudaMemcpyAsync(status, s->d_status, s->p->d.B * sizeof(int), cudaMemcpyDeviceToHost, be->stream));
    CGX_CUDA(cudaStreamSynchronize(be->stream));
  }
  return CGX_OK;
  CGX_CATCH
}

int cgx_solver_try_step(cgx_solver_t s, double alpha, double* dV, int* status) {
  CGX_TRY
  if (!s) throw InvalidArg("Invalid argument: NULL");
  if (alpha > 1. || alpha < 0.)
    throw InvalidArg("Invalid argument: invalid step length, value is between 0. to 1.");  // ddp.cpp:256-259
  Backend* be = s->be;
  start_timing(be);
  be->try_step(alpha);
  finish_timing(be);
  CGX_CUDA(cudaGetLastError());
  if (status) {
    be->export_status(1, s->d_status);
    CGX_CUDA(cudaMemcpyAsync(status, s->d_status, s->p->d.B * sizeof(int), cudaMemcpyDeviceToHost, be->stream));
    CGX_CUDA(cudaStreamSynchronize(be->stream));
  }
  if (dV) {
    fetch_scalars(s);
    for (int b = 0; b < s->p->d.B; ++b) dV[b] = s->h_scal[(size_t)b * 17 + 7];
  }
  return CGX_OK;
  CGX_CATCH
}

int cgx_solver_expected_improvement(cgx_solver_t s, double* d) {
  CGX_TRY
  if (!s || !d) throw InvalidArg("Invalid argument: NULL");
  fetch_scalars(s);
  for (int b = 0; b < s->p->d.B; ++b) {
    d[2 * b] = s->h_scal[(size_t)b * 17 + 2];
    d[2 * b + 1] = s->h_scal[(size_t)b * 17 + 3];
  }
  return CGX_OK;
  CGX_CATCH
}
int cgx_solver_stopping_criteria(cgx_solver_t s, double* stop) {
  CGX_TRY
  if (!s || !stop) throw InvalidArg("Invalid argument: NULL");
  fetch_scalars(s);
  for (int b = 0; b < s->p->d.B; ++b) stop[b] = s->h_scal[(size_t)b * 17 + 1];
  return CGX_OK;
  CGX_CATCH
}

int cgx_solver_get(cgx_solver_t s, const char* name, int b, int t, double* out) {
  CGX_TRY
  if (!s || !name || !out) throw InvalidArg("Invalid argument: NULL");
  FieldInfo fi = s->be->field_info(name);
  if (b < 0 || b >= s->p->d.B || t < 0 || t >= fi.nt) throw InvalidArg("Invalid argument: b / t out of range");
  double* d = scratch(s, (size_t)fi.size * fi.nt);
  s->be->export_field(name, b, 1, d);
  CGX_CUDA(cudaMemcpyAsync(out, d + (size_t)t * fi.size, fi.size * sizeof(double), cudaMemcpyDeviceToHost,
                           s->be->stream));
  CGX_CUDA(cudaStreamSynchronize(s->be->stream));
  CGX_CUDA(cudaGetLastError());
  return fi.size;
  CGX_CATCH
}
int cgx_solver_get_all(cgx_solver_t s, const char* name, double* out) {
  CGX_TRY
  if (!s || !name || !out) throw InvalidArg("Invalid argument: NULL");
  FieldInfo fi = s->be->field_info(name);
  const size_t n = (size_t)s->p->d.B * fi.size * fi.nt;
  double* d = scratch(s, n);
  s->be->export_field(name, 0, s->p->d.B, d);
  CGX_CUDA(cudaMemcpyAsync(out, d, n * sizeof(double), cudaMemcpyDeviceToHost, s->be->stream));
  CGX_CUDA(cudaStreamSynchronize(s->be->stream));
  CGX_CUDA(cudaGetLastError());
  return CGX_OK;
  CGX_CATCH
}
static int set_field(cgx_solver_t s, const char* name, const double* in) {
  CGX_TRY
  if (!s || !in) throw InvalidArg("Invalid argument: NULL");
  FieldInfo fi = s->be->field_info(name);
  const size_t n = (size_t)s->p->d.B * fi.size * fi.nt;
  double* d = scratch(s, n);
  CGX_CUDA(cudaMemcpyAsync(d, in, n * sizeof(double), cudaMemcpyHostToDevice, s->be->stream));
  s->be->import_field(name, d);
  CGX_CUDA(cudaStreamSynchronize(s->be->stream));
  CGX_CUDA(cudaGetLastError());
  return CGX_OK;
  CGX_CATCH
}
int cgx_solver_set_xs(cgx_solver_t s, const double* xs) { return set_field(s, "xs", xs); }
int cgx_solver_set_us(cgx_solver_t s, const double* us) { return set_field(s, "us", us); }

int cgx_solver_get_scalars(cgx_solver_t s, double* out) {
  CGX_TRY
  if (!s || !out) throw InvalidArg("Invalid argument: NULL");
  fetch_scalars(s);
  std::memcpy(out, s->h_scal.data(), s->h_scal.size() * sizeof(double));
  return CGX_OK;
  CGX_CATCH
}

int cgx_solver_get_scalars_row(cgx_solver_t s, int b, double* out) {
  CGX_TRY
  if (!s || !out) throw InvalidArg("Invalid argument: NULL");
  if (b < 0 || b >= s->p->d.B) throw InvalidArg("Invalid argument: b out of range");
  s->be->export_scalars(s->d_scal);
  CGX_CUDA(cudaMemcpyAsync(out, s->d_scal + (size_t)b * 17, 17 * sizeof(double), cudaMemcpyDeviceToHost, s->be->stream));
  CGX_CUDA(cudaStreamSynchronize(s->be->stream));
  return CGX_OK;
  CGX_CATCH
}

int cgx_solver_get_trace(cgx_solver_t s, int b, double* out, int max_records) {
  CGX_TRY
  if (!s) throw InvalidArg("Invalid argument: NULL");
  if (b < 0 || b >= s->p->d.B) throw InvalidArg("Invalid argument: b out of range");
  if (!s->be->trace_on) throw InvalidArg("trace recording is not enabled (cgx_solver_enable_trace)");
  if (out && max_records < 0) throw InvalidArg("Invalid argument: max_records is negative");
  fetch_scalars(s);
  int n = (int)s->h_scal[(size_t)b * 17 + 12];
  if (n > s->be->trace_cap) n = s->be->trace_cap;
  if (!out) return n;
  const int m = n < max_records ? n : max_records;
  double* d = scratch(s, (size_t)s->be->trace_cap * NTRACE);
  s->be->export_trace(b, d);
  CGX_CUDA(cudaMemcpyAsync(out, d, (size_t)m * NTRACE * sizeof(double), cudaMemcpyDeviceToHost, s->be->stream));
  CGX_CUDA(cudaStreamSynchronize(s->be->stream));
  return n;
  CGX_CATCH
}

int cgx_solver_set_callback(cgx_solver_t s, cgx_callback_fn fn, void* user) {
  if (!s) return fail(CGX_ERR_INVALID, "Invalid argument: NULL");
  s->cb = fn;
  s->cb_user = user;
  return CGX_OK;
}

int cgx_solver_last_timing(cgx_solver_t s, double* ms, int* launches) {
  if (!s) return fail(CGX_ERR_INVALID, "Invalid argument: NULL");
  if (ms)
    for (int i = 0; i < 3; ++i) ms[i] = s->be->last_ms[i];
  if (launches) *launches = s->be->last_launches;
  return CGX_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// FP64 DFMA throughput probe (roofline denominator of the Riccati / ABA-derivative kernels)
__global__ void cgx_dfma_probe_kernel(double* out, int iters, double a, double b) {
  double acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += acc[i];
  if (s == 12345.678) out[0] = s;
}

extern "C" int cgx_probe_fp64_tflops(double* tflops) {
  CGX_TRY
  if (!tflops) throw InvalidArg("Invalid argument: NULL");
  require_device();
  double* d = dalloc<double>(1);
  cudaDeviceProp prop;
  int dev = 0;
  CGX_CUDA(cudaGetDevice(&dev));
  CGX_CUDA(cudaGetDeviceProperties(&prop, dev));
  const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 14;
  cudaEvent_t e0, e1;
  CGX_CUDA(cudaEventCreate(&e0));
  CGX_CUDA(cudaEventCreate(&e1));
  double best = 0;
  for (int rep = 0; rep < 5; ++rep) {
    CGX_CUDA(cudaEventRecord(e0));
    cgx_dfma_probe_kernel<<<blocks, threads>>>(d, iters, 0.999999, 1e-9);
    CGX_CUDA(cudaEventRecord(e1));
    CGX_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    CGX_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    const double fl = 2.0 * 8 * (double)iters * blocks * threads;
    const double tf = fl / (ms * 1e-3) / 1e12;
    if (rep > 0 && tf > best) best = tf;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  *tflops = best;
  return CGX_OK;
  CGX_CATCH
}
