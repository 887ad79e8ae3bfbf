// crocoddyl_b200 — cgx_unicycle_calc_noalloc: ShootingProblem::calc for a batch of unicycle problems WITHOUT any device
// allocation. No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// It is a variant of the kernel pair behind cgx_problem_calc on the thread-per-problem backend (tpp_node_data_kernel +
// tpp_cost_sum_kernel, tpp_backend.cu): there the per-node costs go through a cudaMalloc'ed scratch array between the two
// kernels (and the host trajectories through cudaMalloc'ed staging buffers); here all of it lives in ONE file-scope
// static __device__ array of CGX_NOALLOC_SCRATCH_DOUBLES doubles (8 MiB), which every GPU that loads the library holds
// until exit. Same arithmetic, same summation order (shooting.hxx:155-159): results equal cgx_problem_calc's.
//
// The array is shared by every call and keeps the last call's contents. Nothing in it has to start at zero: every word
// a call reads (staged xs / us, per-node costs, totals) is written earlier in the same call, on the same stream. Calls
// that overlap on the GPU (other host threads, other processes' contexts do not share it) RACE on it: serialise them.
#include <mutex>
#include <string>

#include "../../include/crocoddyl_b200.h"
#include "backend.hpp"

namespace cgx {
int set_last_error(int code, const std::string& msg);  // capi.cu: what cgx_last_error() returns
}

namespace {

// In HOST code this name is a host-side stand-in, not a device address: the address a kernel can use is obtained once
// per device with cudaGetSymbolAddress (below). Device code in this file may name the array directly.
static __device__ __align__(256) double cgx_noalloc_scratch[CGX_NOALLOC_SCRATCH_DOUBLES];

constexpr int kMaxDevices = 64;
double* g_scratch_ptr[kMaxDevices] = {};
std::mutex g_scratch_mutex;

struct UniParams {
  double run[3], term[3];
};

// ActionModelUnicycle's cost (unicycle.hxx:38-41), the same operations in the same order as UnicycleModel::cost in
// tpp.cuh (that header defines non-template kernels, so it cannot be included by a second translation unit)
__device__ __forceinline__ double unicycle_cost(const double* p, const double* x, const double* u) {
  const double w0 = p[1], w1 = p[2];
  double rr = 0.;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double r = w0 * x[i];
    rr += r * r;
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const double r = w1 * u[i];
    rr += r * r;
  }
  return 0.5 * rr;
}

// per-node cost of problem b, node t (tpp_node_data_kernel's calc branch: unicycle.hxx:38-41; the terminal node is the
// terminal model evaluated with u = unone = 0, action-base.hxx:27-31)
__global__ void noalloc_unicycle_node_cost_kernel(const double* __restrict__ xs, const double* __restrict__ us, int B, int T,
                                                  UniParams P, double* __restrict__ node_cost) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  const int t = blockIdx.y;
  if (b >= B) return;
  const size_t node = (size_t)b * (T + 1) + t;
  double x[3], u[2];
#pragma unroll
  for (int i = 0; i < 3; ++i) x[i] = xs[node * 3 + i];
#pragma unroll
  for (int i = 0; i < 2; ++i) u[i] = (t < T) ? us[((size_t)b * T + (t < T ? t : 0)) * 2 + i] : 0.;
  node_cost[node] = unicycle_cost(t < T ? P.run : P.term, x, u);
}
// total cost in the reference's order: running nodes t = 0..T-1, then the terminal node (shooting.hxx:155-159)
__global__ void noalloc_cost_sum_kernel(int B, int T, const double* __restrict__ node_cost, double* __restrict__ costs) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double s = 0.;
  for (int t = 0; t <= T; ++t) s += node_cost[(size_t)b * (T + 1) + t];
  costs[b] = s;
}

// device address of the static scratch array on the CURRENT device (looked up once per device, with it current)
double* scratch_on_current_device() {
  int dev = 0;
  CGX_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) throw cgx::Unsupported("cgx_unicycle_calc_noalloc: device index beyond 63");
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_ptr[dev]) {
    double* ptr = nullptr;
    CGX_CUDA(cudaGetSymbolAddress((void**)&ptr, cgx_noalloc_scratch));
    if (!ptr) throw cgx::CudaError("cudaGetSymbolAddress returned a null address for the no-alloc scratch array");
    g_scratch_ptr[dev] = ptr;
  }
  return g_scratch_ptr[dev];
}

}  // namespace

extern "C" int cgx_unicycle_calc_noalloc(int B, int T, const double* params, const double* tparams, const double* xs,
                                         const double* us, double* costs) {
  try {
    if (B <= 0 || T <= 0) throw cgx::InvalidArg("Invalid argument: B and T must be positive");
    if (!params || !xs || !us || !costs) throw cgx::InvalidArg("Invalid argument: params / xs / us / costs is NULL");
    const size_t n_xs = (size_t)B * (T + 1) * 3, n_us = (size_t)B * T * 2, n_nc = (size_t)B * (T + 1), n_c = (size_t)B;
    const size_t need = n_xs + n_us + n_nc + n_c;
    if (need > (size_t)CGX_NOALLOC_SCRATCH_DOUBLES)
      throw cgx::InvalidArg("Invalid argument: cgx_unicycle_calc_noalloc needs B * (6 T + 5) = " + std::to_string(need) +
                            " doubles of scratch, the static array holds CGX_NOALLOC_SCRATCH_DOUBLES = " +
                            std::to_string((size_t)CGX_NOALLOC_SCRATCH_DOUBLES) + " (no allocation is made: use cgx_problem_calc "
                            "for larger batches)");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
      cudaGetLastError();
      throw cgx::CudaError("no CUDA device available: crocoddyl_b200 has no CPU fallback");
    }
    double* scratch = scratch_on_current_device();
    double* d_xs = scratch;
    double* d_us = d_xs + n_xs;
    double* d_nc = d_us + n_us;
    double* d_c = d_nc + n_nc;
    UniParams P;
    for (int i = 0; i < 3; ++i) {
      P.run[i] = params[i];
      P.term[i] = tparams ? tparams[i] : params[i];
    }
    // cudaMemcpyDefault: xs / us / costs may be host or device pointers (unified addressing); legacy default stream,
    // no stream, event or memory is created
    CGX_CUDA(cudaMemcpyAsync(d_xs, xs, n_xs * sizeof(double), cudaMemcpyDefault, 0));
    CGX_CUDA(cudaMemcpyAsync(d_us, us, n_us * sizeof(double), cudaMemcpyDefault, 0));
    noalloc_unicycle_node_cost_kernel<<<dim3((B + 63) / 64, T + 1), 64, 0, 0>>>(d_xs, d_us, B, T, P, d_nc);
    noalloc_cost_sum_kernel<<<(B + 63) / 64, 64, 0, 0>>>(B, T, d_nc, d_c);
    CGX_CUDA(cudaGetLastError());
    CGX_CUDA(cudaMemcpyAsync(costs, d_c, n_c * sizeof(double), cudaMemcpyDefault, 0));
    CGX_CUDA(cudaStreamSynchronize(0));
    return CGX_OK;
  } catch (const cgx::CudaError& e) {
    return cgx::set_last_error(CGX_ERR_CUDA, e.what());
  } catch (const cgx::InvalidArg& e) {
    return cgx::set_last_error(CGX_ERR_INVALID, e.what());
  } catch (const cgx::Unsupported& e) {
    return cgx::set_last_error(CGX_ERR_UNSUPPORTED, e.what());
  } catch (const std::exception& e) {
    return cgx::set_last_error(CGX_ERR_INVALID, e.what());
  }
}
