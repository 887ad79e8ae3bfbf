// crocoddyl_b200 — host driver of the thread-per-problem kernels (tpp.cuh). Model family: unicycle.
#include <cstring>
#include <map>
#include <cstdlib>
#include "backend.hpp"
#include "tpp.cuh"

namespace cgx {

namespace {

// launch kernel<M, BOX> with BOX = c.box: the control-limited solvers are separate instantiations, so the register
// budget and the code of the unconstrained solve kernel are untouched by them
#define CGX_UNPAREN(...) __VA_ARGS__
#define TPP_LAUNCH_BOX(kernel, cfg, ...)                        \
  do {                                                          \
    if (c.box)                                                  \
      kernel<M, true><<<CGX_UNPAREN cfg>>>(__VA_ARGS__);        \
    else                                                        \
      kernel<M, false><<<CGX_UNPAREN cfg>>>(__VA_ARGS__);       \
  } while (0)

template <class M>
struct TppBackend : Backend {
  static constexpr int N = M::NX, U = M::NU;
  TppCtx c;
  std::vector<void*> allocs;
  bool solver_allocated = false;
  int* d_list = nullptr;  // compaction: [Bp] problem indices, per-block counts, number of unfinished problems
  double* node_data[10] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};

  explicit TppBackend(const ProblemDesc* p_, int sk) {
    p = p_;
    solver_kind = sk;
    std::memset(&c, 0, sizeof(c));
    c.B = p->B;
    c.Bp = ((p->B + 63) / 64) * 64;
    c.T = p->T;
    c.fddp = sk == 1 || sk == 3;  // SolverFDDP / SolverBoxFDDP
    c.box = sk >= 2;              // SolverBoxDDP / SolverBoxFDDP
    CGX_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    for (int i = 0; i < 4; ++i) CGX_CUDA(cudaEventCreate(&ev[i]));
    c.mparams = p->d_params;
    c.tparams = p->d_tparams;
    c.u_lb = p->d_ulb;
    c.u_ub = p->d_uub;
    c.has_limits = p->d_haslim;
    c.x0 = alloc((size_t)N * c.Bp);
    // 8 resident blocks of the solve kernel per SM (the whole batch in one wave) need 8 x 24.6 KB of staging area
    CGX_CUDA(cudaFuncSetAttribute(tpp_solve_kernel<M, false>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  cudaSharedmemCarveoutMaxShared));
    CGX_CUDA(cudaFuncSetAttribute(tpp_solve_kernel<M, true>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  cudaSharedmemCarveoutMaxShared));
    x0_changed();
  }
  ~TppBackend() override {
    for (void* q : allocs) cudaFree(q);
    for (int i = 0; i < 4; ++i)
      if (ev[i]) cudaEventDestroy(ev[i]);
    for (int k = 0; k < PIPE_MAX; ++k) {
      if (pev[k]) cudaEventDestroy(pev[k]);
      for (int j = 0; j < 3; ++j)
        if (ptime[k][j]) cudaEventDestroy(ptime[k][j]);
      if (pstream[k]) cudaStreamDestroy(pstream[k]);
    }
    if (ev_reset) cudaEventDestroy(ev_reset);
    if (stream) cudaStreamDestroy(stream);
  }
  double* alloc(size_t n) {
    double* q = dalloc<double>(n);
    CGX_CUDA(cudaMemsetAsync(q, 0, n * sizeof(double), stream));
    allocs.push_back(q);
    return q;
  }
  void release(double*& q) {
    if (!q) return;
    for (size_t i = 0; i < allocs.size(); ++i)
      if (allocs[i] == q) {
        allocs.erase(allocs.begin() + i);
        break;
      }
    cudaFree(q);
    q = nullptr;
  }
  typedef TppLayout<M> L;
  void gather(double* d_user, const TppFieldMap& fm, double* b0p, double* b1p, const int* sel, int flip, int b0,
              int nb, int to_rec, cudaStream_t st = nullptr) {
    tpp_gather_kernel<<<dim3((nb + 63) / 64, fm.nt), 64, 0, st ? st : stream>>>(d_user, b0p, b1p, sel, flip, fm, c.Bp,
                                                                                 b0, nb, to_rec);
    ++last_launches;
  }
  int nblocks() const { return (c.B + 63) / 64; }
  const int* cur_sel() const { return c.iscal + (size_t)I_CUR * c.Bp; }
  const int* fs_sel() const { return c.iscal + (size_t)I_FSBUF * c.Bp; }

  void x0_changed() override {
    tpp_x0_kernel<<<nblocks(), 64, 0, stream>>>(p->d_x0, c.x0, N, c.B, c.Bp);
    CGX_CUDA(cudaStreamSynchronize(stream));
  }

  void configure() override {
    c.o = o;
    const size_t Bp = c.Bp, T = c.T;
    if (!solver_allocated) {
      for (int i = 0; i < 2; ++i) c.xu[i] = alloc((T + 1) * Bp * L::R_XU);
      c.kk = alloc(T * Bp * L::R_KK);
      c.fs = alloc((T + 1) * Bp * L::R_FS);
      c.w = alloc((T + 1) * Bp * L::R_FS);
      c.vx = alloc((T + 1) * Bp * L::R_FS);
      c.vl = alloc((T + 1) * Bp * L::R_VL);
      c.scal = alloc((size_t)S_NSLOTS * Bp);
      int* is = dalloc<int>((size_t)I_NSLOTS * Bp);
      allocs.push_back(is);
      c.iscal = is;
      CGX_CUDA(cudaMemsetAsync(is, 0, (size_t)I_NSLOTS * Bp * sizeof(int), stream));
      // SolverAbstract ctor values (solver-base.cpp:14-25): xreg = ureg = NaN, steplength = 1
      std::vector<double> h((size_t)S_NSLOTS * Bp, 0.);
      for (size_t b = 0; b < Bp; ++b) {
        h[(size_t)S_XREG * Bp + b] = NAN;
        h[(size_t)S_UREG * Bp + b] = NAN;
        h[(size_t)S_STEP * Bp + b] = 1.;
      }
      CGX_CUDA(cudaMemcpyAsync(c.scal, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice, stream));
      CGX_CUDA(cudaStreamSynchronize(stream));
      solver_allocated = true;
    }
    if (full && !c.q) c.q = alloc(T * Bp * L::R_Q);
    if (c.box && !c.quu_inv) c.quu_inv = alloc(T * Bp * U * U);
    c.full = full;
    // lean storage: Vx / Vxx are materialised once per solve (tpp.cuh); full storage and the box solvers (whose QP is
    // warm-started from the previous gains) store them in every backward pass
    c.store_v = (full || c.box) ? 1 : 0;
    if (trace_on && (!c.trace || c.trace_cap != trace_cap)) {
      release(c.trace);
      c.trace = alloc((size_t)trace_cap * NTRACE * Bp);
      c.trace_cap = trace_cap;
    }
    if (!trace_on) {
      release(c.trace);
      c.trace_cap = 0;
    }
  }

  void set_candidate(const double* d_xs, const double* d_us, int is_feasible) override {
    const size_t Bp = c.Bp, T = c.T;
    tpp_reset_kernel<<<(c.Bp + 63) / 64, 64, 0, stream>>>(c, is_feasible);
    ++last_launches;
    if (!d_xs || !d_us) CGX_CUDA(cudaMemsetAsync(c.xu[0], 0, (T + 1) * Bp * L::R_XU * sizeof(double), stream));
    if (d_xs) gather(const_cast<double*>(d_xs), fmap("xs"), c.xu[0], c.xu[0], nullptr, 0, 0, c.B, 1);
    if (d_us) gather(const_cast<double*>(d_us), fmap("us"), c.xu[0], c.xu[0], nullptr, 0, 0, c.B, 1);
  }

  // Optional chunked solve: `chunk` iterations per launch, unfinished problems compacted in between (tpp.cuh).
  // MEASURED on B200 (config C2, B = 65,536, profiles/README.md): one launch 51.1 ms; chunk 8 / 4 / 2 = 53.9 / 58.0 /
  // 65.5 ms. The loop is bound by the dependent instruction chain of each OCP (0.3 ms per iteration for a lone warp)
  // and by instruction issue while all OCPs are alive, not by idle lanes, and compaction scatters the lanes' records
  // over DRAM pages. So the default stays ONE launch (0); CROCODDYL_B200_TPP_CHUNK=<n> enables the chunked path.
  // iteration after which the unfinished problems of a large batch are compacted and handed to the tail kernel
  // (0 = never). CROCODDYL_B200_TPP_TAIL_AT overrides (development knob).
  // Measured on B200 (C2, B = 65,536): the resident solve is fastest with the cut at 18-20 (24.2-24.7 ms against 24.9 at
  // 22 and 25.7 at 30); the sliced host pipeline, whose slices of 16,384 problems leave the machine emptier, prefers a
  // later cut (e2e 34.4-35.3 ms at 26-28 against 35.2-36.2 at 22).
  int tail_iter(bool pipelined = false) const {
    const char* e = getenv("CROCODDYL_B200_TPP_TAIL_AT");
    return e ? atoi(e) : (pipelined ? 27 : 19);
  }
  // compaction scratch: [Bp] problem indices, per-1024-block counts (room for PIPE_MAX ragged slices), counters
  void ensure_list() {
    if (d_list) return;
    d_list = dalloc<int>((size_t)c.Bp + (c.B + 1023) / 1024 + 2 * PIPE_MAX + 2);
    allocs.push_back(d_list);
  }
  int chunk_iters() const {
    const char* e = getenv("CROCODDYL_B200_TPP_CHUNK");
    return e ? atoi(e) : 0;
  }
  // solve() of a batch that arrives in HOST buffers and whose xs/us go back to HOST buffers: the batch is cut into K
  // slices, each with its own stream: H2D copy of the slice -> layout conversion -> the whole solve loop of the slice
  // -> conversion back -> D2H copy. The PCIe copies of one slice overlap the solve of the others (all slices are
  // co-resident: 65,536 OCPs are one wave), the arithmetic per OCP is the same kernel as solve().
  // CROCODDYL_B200_TPP_PIPELINE=<K> (default 4; 0/1 = off, the caller then uses upload / solve / get_all in sequence).
  static constexpr int PIPE_MAX = 8;
  cudaStream_t pstream[PIPE_MAX] = {};
  cudaStream_t dstream[PIPE_MAX] = {};  // early download of a slice while its tail kernel runs
  cudaEvent_t pev[PIPE_MAX] = {};
  cudaEvent_t pev_phase1[PIPE_MAX] = {}, pev_early[PIPE_MAX] = {};
  cudaEvent_t ev_reset = nullptr;
  cudaEvent_t ptime[PIPE_MAX][3] = {};  // CROCODDYL_B200_TPP_PIPELINE_TRACE=1: upload done / solve done / download done
  double* d_out_xs = nullptr;
  double* d_out_us = nullptr;
  int pipeline_chunks() const {
    const char* e = getenv("CROCODDYL_B200_TPP_PIPELINE");
    const int k = e ? atoi(e) : 4;
    return k > PIPE_MAX ? PIPE_MAX : k;
  }
  // device-usable address of a page-locked host buffer (unified addressing), or null if p is pageable / device memory
  static double* mapped_host(double* p) {
    if (!p) return nullptr;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
      cudaGetLastError();
      return nullptr;
    }
    return at.type == cudaMemoryTypeHost ? (double*)at.devicePointer : nullptr;
  }
  bool solve_pipelined(const double* h_xs, const double* h_us, double* d_xs, double* d_us, int maxiter,
                       int is_feasible, double reginit, double* h_xs_out, double* h_us_out) override {
    const int K = pipeline_chunks();
    if (K < 2 || c.B < 1024 * K || chunk_iters() > 0 || !h_xs || !h_us) return false;
    c.o = o;
    const size_t T = c.T;
    if (!ev_reset) CGX_CUDA(cudaEventCreateWithFlags(&ev_reset, cudaEventDisableTiming));
    for (int k = 0; k < K; ++k)
      if (!pstream[k]) {
        CGX_CUDA(cudaStreamCreateWithFlags(&pstream[k], cudaStreamNonBlocking));
        CGX_CUDA(cudaEventCreateWithFlags(&pev[k], cudaEventDisableTiming));
      }
    const char* pte = getenv("CROCODDYL_B200_TPP_PIPELINE_TRACE");
    const bool ptrace = pte && atoi(pte) != 0;
    if (ptrace)
      for (int k = 0; k < K; ++k)
        for (int j = 0; j < 3; ++j)
          if (!ptime[k][j]) CGX_CUDA(cudaEventCreate(&ptime[k][j]));
    if (h_xs_out && !d_out_xs) d_out_xs = alloc((size_t)c.B * (T + 1) * N);
    if (h_us_out && !d_out_us) d_out_us = alloc((size_t)c.B * T * U);
    tpp_reset_kernel<<<(c.Bp + 63) / 64, 64, 0, stream>>>(c, is_feasible);
    ++last_launches;
    CGX_CUDA(cudaEventRecord(ev_reset, stream));
    CGX_CUDA(cudaEventRecord(ev[1], stream));
    const int per = ((c.B + K - 1) / K + 63) / 64 * 64;
    const Fld fx = fld("xs"), fu = fld("us");
    const int tail_at = tail_iter(true);
    const bool two_phase = tail_at > 0 && maxiter > tail_at && c.B >= 4096;
    if (two_phase) ensure_list();
    // Early download (two-phase solve into page-locked output buffers): the problems that finished in phase 1 — the
    // majority — are final when the tail kernel starts, so the WHOLE slice is converted and copied out on a second
    // stream while the tail kernel runs; afterwards only the tail's problems are written again, straight into the
    // caller's buffers through the unified address space (coalesced per problem). The early copy of an unfinished
    // problem holds an intermediate iterate and is overwritten; the order is enforced by pev_early.
    // CROCODDYL_B200_TPP_EARLY_DOWNLOAD=0 disables it (the copies back then follow the tail kernel).
    const char* ede = getenv("CROCODDYL_B200_TPP_EARLY_DOWNLOAD");
    double* m_xs_out = mapped_host(h_xs_out);
    double* m_us_out = mapped_host(h_us_out);
    const bool early = two_phase && !(ede && atoi(ede) == 0) && (h_xs_out || h_us_out) && (!h_xs_out || m_xs_out) &&
                       (!h_us_out || m_us_out);
    if (early)
      for (int k = 0; k < K; ++k)
        if (!dstream[k]) {
          CGX_CUDA(cudaStreamCreateWithFlags(&dstream[k], cudaStreamNonBlocking));
          CGX_CUDA(cudaEventCreateWithFlags(&pev_phase1[k], cudaEventDisableTiming));
          CGX_CUDA(cudaEventCreateWithFlags(&pev_early[k], cudaEventDisableTiming));
        }
    int cb0 = 0;
    for (int k = 0; k < K; ++k) {
      const int b0 = k * per;
      if (b0 >= c.B) break;
      const int nb = c.B - b0 < per ? c.B - b0 : per;
      cudaStream_t st = pstream[k];
      const size_t ox = (size_t)b0 * (T + 1) * N, nx = (size_t)nb * (T + 1) * N * sizeof(double);
      const size_t ou = (size_t)b0 * T * U, nu = (size_t)nb * T * U * sizeof(double);
      CGX_CUDA(cudaStreamWaitEvent(st, ev_reset, 0));
      CGX_CUDA(cudaMemcpyAsync(d_xs + ox, h_xs + ox, nx, cudaMemcpyHostToDevice, st));
      CGX_CUDA(cudaMemcpyAsync(d_us + ou, h_us + ou, nu, cudaMemcpyHostToDevice, st));
      gather(d_xs + ox, fx.fm, c.xu[0], c.xu[0], nullptr, 0, b0, nb, 1, st);
      gather(d_us + ou, fu.fm, c.xu[0], c.xu[0], nullptr, 0, b0, nb, 1, st);
      if (ptrace) CGX_CUDA(cudaEventRecord(ptime[k][0], st));
      if (two_phase) {  // as solve(): the slice's unfinished problems are compacted and finished by the tail kernel
        const int cb = (nb + 1023) / 1024;
        int* cnt = d_list + c.Bp + cb0;
        int* nact = d_list + c.Bp + (c.B + 1023) / 1024 + PIPE_MAX + 1 + k;
        cb0 += cb;
        const int* fin = c.iscal + (size_t)I_FINISHED * c.Bp + b0;
        TPP_LAUNCH_BOX(tpp_solve_kernel, ((nb + 63) / 64, TPP_BLOCK, 0, st), c, maxiter, is_feasible, reginit, tail_at, 1, nullptr,
                       nullptr, b0, b0 + nb);
        tpp_compact_count_kernel<<<cb, 1024, 0, st>>>(fin, nb, cnt);
        tpp_compact_scatter_kernel<<<cb, 1024, 0, st>>>(fin, nb, cnt, d_list + b0, nact, b0);
        if (early) CGX_CUDA(cudaEventRecord(pev_phase1[k], st));
        TPP_LAUNCH_BOX(tpp_solve_tail_kernel, ((nb + 31) / 32, 32, 0, st), c, maxiter, is_feasible, reginit, 0, 0, d_list + b0, nact, 0,
                       nb);
        last_launches += 4;
      } else {
        TPP_LAUNCH_BOX(tpp_solve_kernel, ((nb + 63) / 64, TPP_BLOCK, 0, st), c, maxiter, is_feasible, reginit, 0, 1, nullptr, nullptr,
                       b0, b0 + nb);
        ++last_launches;
      }
      if (ptrace) CGX_CUDA(cudaEventRecord(ptime[k][1], st));
    }
    // second pass in issue order: the copies back are enqueued after EVERY slice's upload and solve, so that no
    // upload can queue up behind a download that waits for a solve
    for (int k = 0; k < K; ++k) {
      const int b0 = k * per;
      if (b0 >= c.B) break;
      const int nb = c.B - b0 < per ? c.B - b0 : per;
      cudaStream_t st = pstream[k];
      const size_t ox = (size_t)b0 * (T + 1) * N, nx = (size_t)nb * (T + 1) * N * sizeof(double);
      const size_t ou = (size_t)b0 * T * U, nu = (size_t)nb * T * U * sizeof(double);
      if (early) {
        cudaStream_t ds = dstream[k];
        CGX_CUDA(cudaStreamWaitEvent(ds, pev_phase1[k], 0));
        if (h_xs_out) {
          gather(d_out_xs + ox, fx.fm, fx.a, fx.b, fx.sel, fx.flip, b0, nb, 0, ds);
          CGX_CUDA(cudaMemcpyAsync(h_xs_out + ox, d_out_xs + ox, nx, cudaMemcpyDeviceToHost, ds));
        }
        if (h_us_out) {
          gather(d_out_us + ou, fu.fm, fu.a, fu.b, fu.sel, fu.flip, b0, nb, 0, ds);
          CGX_CUDA(cudaMemcpyAsync(h_us_out + ou, d_out_us + ou, nu, cudaMemcpyDeviceToHost, ds));
        }
        CGX_CUDA(cudaEventRecord(pev_early[k], ds));
        // the tail's problems, after the tail kernel AND after the early copy of this slice has landed
        CGX_CUDA(cudaStreamWaitEvent(st, pev_early[k], 0));
        const int gl = nb < 4096 ? nb : 4096;
        if (h_xs_out) {
          tpp_gather_list_kernel<<<gl, 128, 0, st>>>(m_xs_out, fx.a, fx.b, fx.sel, fx.flip, fx.fm, c.Bp, d_list + b0, d_list + c.Bp + (c.B + 1023) / 1024 + PIPE_MAX + 1 + k);
          ++last_launches;
        }
        if (h_us_out) {
          tpp_gather_list_kernel<<<gl, 128, 0, st>>>(m_us_out, fu.a, fu.b, fu.sel, fu.flip, fu.fm, c.Bp, d_list + b0, d_list + c.Bp + (c.B + 1023) / 1024 + PIPE_MAX + 1 + k);
          ++last_launches;
        }
      }
      if (h_xs_out && !early) {
        gather(d_out_xs + ox, fx.fm, fx.a, fx.b, fx.sel, fx.flip, b0, nb, 0, st);
        CGX_CUDA(cudaMemcpyAsync(h_xs_out + ox, d_out_xs + ox, nx, cudaMemcpyDeviceToHost, st));
      }
      if (h_us_out && !early) {
        gather(d_out_us + ou, fu.fm, fu.a, fu.b, fu.sel, fu.flip, b0, nb, 0, st);
        CGX_CUDA(cudaMemcpyAsync(h_us_out + ou, d_out_us + ou, nu, cudaMemcpyDeviceToHost, st));
      }
      if (ptrace) CGX_CUDA(cudaEventRecord(ptime[k][2], st));
      CGX_CUDA(cudaEventRecord(pev[k], st));
      CGX_CUDA(cudaStreamWaitEvent(stream, pev[k], 0));
    }
    CGX_CUDA(cudaEventRecord(ev[2], stream));
    if (ptrace) {  // development aid: device timeline of the slices relative to the start of the call (ms)
      CGX_CUDA(cudaStreamSynchronize(stream));
      for (int k = 0; k < K && k * per < c.B; ++k) {
        float a = 0, b = 0, d = 0;
        cudaEventElapsedTime(&a, ev[1], ptime[k][0]);
        cudaEventElapsedTime(&b, ev[1], ptime[k][1]);
        cudaEventElapsedTime(&d, ev[1], ptime[k][2]);
        fprintf(stderr, "[tpp pipeline] slice %d: upload done %.2f, solve done %.2f, download done %.2f ms\n", k, a, b, d);
      }
    }
    CGX_CUDA(cudaGetLastError());
    return true;
  }

  void solve(int maxiter, int is_feasible, double reginit) override {
    c.o = o;
    const int chunk = chunk_iters();
    const int tail_at = tail_iter();
    CGX_CUDA(cudaEventRecord(ev[1], stream));
    auto compaction = [&](int*& d_counts, int*& d_nactive, int& cblocks) {
      cblocks = (c.B + 1023) / 1024;
      ensure_list();
      d_counts = d_list + c.Bp;
      d_nactive = d_counts + cblocks + PIPE_MAX;
    };
    if (chunk > 0 && maxiter > chunk) {
      int *d_counts, *d_nactive, cblocks;
      compaction(d_counts, d_nactive, cblocks);
      const int* fin = c.iscal + (size_t)I_FINISHED * c.Bp;
      TPP_LAUNCH_BOX(tpp_solve_kernel, (nblocks(), TPP_BLOCK, 0, stream), c, maxiter, is_feasible, reginit, chunk, 1, nullptr, nullptr, 0, c.B);
      ++last_launches;
      for (int done = chunk; done < maxiter; done += chunk) {
        tpp_compact_count_kernel<<<cblocks, 1024, 0, stream>>>(fin, c.B, d_counts);
        tpp_compact_scatter_kernel<<<cblocks, 1024, 0, stream>>>(fin, c.B, d_counts, d_list, d_nactive, 0);
        TPP_LAUNCH_BOX(tpp_solve_kernel, (nblocks(), TPP_BLOCK, 0, stream), c, maxiter, is_feasible, reginit, chunk, 0, d_list, d_nactive, 0, c.B);
        last_launches += 3;
      }
    } else if (tail_at > 0 && maxiter > tail_at && c.B >= 4096) {
      // Two regimes of a large batch: while most problems iterate the launch is bound by HBM traffic and
      // instruction issue at full occupancy; once most have returned, what is left is the dependent chain of the
      // slowest problems. After `tail_at` iterations the unfinished problems are compacted (ordered) and finished by
      // the tail instantiation: full warps again, one warp per block, no register cap. Same device functions, same
      // per-problem arithmetic: results are bit-identical to the single launch.
      int *d_counts, *d_nactive, cblocks;
      compaction(d_counts, d_nactive, cblocks);
      const int* fin = c.iscal + (size_t)I_FINISHED * c.Bp;
      TPP_LAUNCH_BOX(tpp_solve_kernel, (nblocks(), TPP_BLOCK, 0, stream), c, maxiter, is_feasible, reginit, tail_at, 1, nullptr, nullptr, 0, c.B);
      tpp_compact_count_kernel<<<cblocks, 1024, 0, stream>>>(fin, c.B, d_counts);
      tpp_compact_scatter_kernel<<<cblocks, 1024, 0, stream>>>(fin, c.B, d_counts, d_list, d_nactive, 0);
      TPP_LAUNCH_BOX(tpp_solve_tail_kernel, ((c.B + 31) / 32, 32, 0, stream), c, maxiter, is_feasible, reginit, 0, 0, d_list, d_nactive, 0, c.B);
      last_launches += 4;
    } else {
      TPP_LAUNCH_BOX(tpp_solve_kernel, (nblocks(), TPP_BLOCK, 0, stream), c, maxiter, is_feasible, reginit, 0, 1, nullptr, nullptr, 0, c.B);
      ++last_launches;
    }
    CGX_CUDA(cudaEventRecord(ev[2], stream));
    CGX_CUDA(cudaGetLastError());
  }
  void compute_direction(int recalc) override {
    c.o = o;
    CGX_CUDA(cudaEventRecord(ev[1], stream));
    TPP_LAUNCH_BOX(tpp_direction_kernel, (nblocks(), 64, 0, stream), c, recalc);
    CGX_CUDA(cudaEventRecord(ev[2], stream));
    ++last_launches;
    CGX_CUDA(cudaGetLastError());
  }
  void try_step(double alpha) override {
    c.o = o;
    CGX_CUDA(cudaEventRecord(ev[1], stream));
    TPP_LAUNCH_BOX(tpp_trystep_kernel, (nblocks(), 64, 0, stream), c, alpha);
    CGX_CUDA(cudaEventRecord(ev[2], stream));
    ++last_launches;
    CGX_CUDA(cudaGetLastError());
  }

  struct Fld {
    TppFieldMap fm;
    double* a;
    double* b;
    const int* sel;
    int flip;
  };
  TppFieldMap fmap(const std::string& n) { return fld(n).fm; }
  Fld fld(const std::string& n) {
    const int T = c.T;
    Fld f;
    f.sel = nullptr;
    f.flip = 0;
    auto seq = [&](int R, int size, int nt, int off0) {
      f.fm.R = R;
      f.fm.size = size;
      f.fm.nt = nt;
      if (size > 16) throw Unsupported("field too wide for the thread-per-problem backend");
      for (int i = 0; i < size; ++i) f.fm.off[i] = off0 + i;
    };
    if (n == "xs" || n == "xs_try" || n == "us" || n == "us_try") {
      f.a = c.xu[0];
      f.b = c.xu[1];
      f.sel = cur_sel();
      f.flip = (n == "xs_try" || n == "us_try") ? 1 : 0;
      if (n[0] == 'x') seq(L::R_XU, N, T + 1, L::XU_X);
      else seq(L::R_XU, U, T, L::XU_U);
      return f;
    }
    if (n == "fs") {
      f.a = f.b = c.fs;
      seq(L::R_FS, N, T + 1, 0);
      return f;
    }
    if (n == "K" || n == "k") {
      f.a = f.b = c.kk;
      if (n == "K") seq(L::R_KK, U * N, T, L::KK_K);
      else seq(L::R_KK, U, T, L::KK_k);
      return f;
    }
    if (n == "Quu_inv") {  // SolverBoxDDP::get_Quu_inv (box-ddp.hpp)
      if (!c.quu_inv) throw Unsupported("Quu_inv is kept by the box solvers only");
      f.a = f.b = c.quu_inv;
      seq(U * U, U * U, T, 0);
      return f;
    }
    if (n == "Vx") {
      f.a = f.b = c.vx;
      seq(L::R_FS, N, T + 1, 0);
      return f;
    }
    if (n == "Vxx") {
      f.a = f.b = c.vl;
      seq(L::R_VL, N * N, T + 1, 0);
      for (int j = 0; j < N; ++j)
        for (int i = 0; i < N; ++i) f.fm.off[i + N * j] = L::sym(i, j);
      return f;
    }
    if (n == "Qxx" || n == "Qxu" || n == "Quu" || n == "Qx" || n == "Qu") {
      if (!c.full || !c.q) throw Unsupported("field " + n + " is not kept with cgx_solver_set_storage(full = 0)");
      f.a = f.b = c.q;
      if (n == "Qxx") seq(L::R_Q, N * N, T, L::Q_XX);
      else if (n == "Qxu") seq(L::R_Q, N * U, T, L::Q_XU);
      else if (n == "Quu") seq(L::R_Q, U * U, T, L::Q_UU);
      else if (n == "Qx") seq(L::R_Q, N, T, L::Q_X);
      else seq(L::R_Q, U, T, L::Q_U);
      return f;
    }
    throw NotFound("unknown field " + n);
  }
  FieldInfo field_info(const std::string& name) override {
    Fld f = fld(name);
    return {f.fm.size, f.fm.nt};
  }
  void export_field(const std::string& name, int b0, int nb, double* d_out) override {
    Fld f = fld(name);
    gather(d_out, f.fm, f.a, f.b, f.sel, f.flip, b0, nb, 0);
  }
  void import_field(const std::string& name, const double* d_in) override {
    if (name != "xs" && name != "us") throw Unsupported("only xs and us can be set");
    Fld f = fld(name);
    gather(const_cast<double*>(d_in), f.fm, f.a, f.b, f.sel, f.flip, 0, c.B, 1);
  }
  void export_scalars(double* d_out) override;
  void export_status(int which, int* d_out) override {
    CGX_CUDA(cudaMemcpyAsync(d_out, c.iscal + (size_t)(which ? I_FSTATUS : I_BSTATUS) * c.Bp, c.B * sizeof(int),
                             cudaMemcpyDeviceToDevice, stream));
  }
  void set_reg(double xreg, double ureg, bool set_x, bool set_u) override;
  void export_trace(int b, double* d_out) override;
  void export_trace_all(int nrec, double* d_out) override;

  // ---- problem-level API ----
  void problem_calc(const double* d_xs, const double* d_us, int diff, double* d_costs) override;
  void problem_rollout(const double* d_us, double* d_xs) override {
    tpp_rollout_kernel<M><<<nblocks(), 64, 0, stream>>>(c, d_us, d_xs);
    ++last_launches;
  }
  void problem_quasi_static(const double*, double*) override {
    throw Unsupported("quasiStatic is not implemented for this model family");
  }
  FieldInfo data_info(const std::string& name) override {
    const int T = c.T;
    if (name == "cost") return {1, T + 1};
    if (name == "xnext") return {N, T + 1};
    if (name == "Fx") return {N * N, T + 1};
    if (name == "Fu") return {N * U, T + 1};
    if (name == "Lx") return {N, T + 1};
    if (name == "Lu") return {U, T + 1};
    if (name == "Lxx") return {N * N, T + 1};
    if (name == "Lxu") return {N * U, T + 1};
    if (name == "Luu") return {U * U, T + 1};
    if (name == "r") return {M::NR, T + 1};  // ActionDataAbstract::r (action-base.hpp:130), written by calc
    throw NotFound("unknown data field " + name);
  }
  void export_data(const std::string& name, int b, int t, double* d_out) override {
    static const char* names[10] = {"cost", "xnext", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu", "r"};
    FieldInfo fi = data_info(name);
    for (int i = 0; i < 10; ++i)
      if (name == names[i]) {
        if (!node_data[i]) throw InvalidArg("node data not computed yet: call calc / calcDiff first");
        CGX_CUDA(cudaMemcpyAsync(d_out, node_data[i] + ((size_t)b * (c.T + 1) + t) * fi.size,
                                 fi.size * sizeof(double), cudaMemcpyDeviceToDevice, stream));
      }
  }
};

__global__ void tpp_scalars_kernel(TppCtx c, double* out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  const int Bp = c.Bp;
  double* o = out + (size_t)b * 17;
  o[0] = TPP_S(S_COST);
  o[1] = TPP_S(S_STOP);
  o[2] = TPP_S(S_D0);
  o[3] = TPP_S(S_D1);
  o[4] = TPP_S(S_XREG);
  o[5] = TPP_S(S_UREG);
  o[6] = TPP_S(S_STEP);
  o[7] = TPP_S(S_DV);
  o[8] = TPP_S(S_DVEXP);
  o[9] = TPP_I(I_ITER);
  o[10] = TPP_I(I_FEAS);
  o[11] = TPP_I(I_CONVERGED);
  o[12] = TPP_I(I_NREC);
  o[13] = TPP_S(S_COST_TRY);
  o[14] = TPP_I(I_WASFEAS);
  o[15] = TPP_I(I_NBWD);
  o[16] = TPP_I(I_NFWD);
}
__global__ void tpp_setreg_kernel(TppCtx c, double xreg, double ureg, int set_x, int set_u) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  const int Bp = c.Bp;
  if (set_x) TPP_S(S_XREG) = xreg;
  if (set_u) TPP_S(S_UREG) = ureg;
}
__global__ void tpp_trace_kernel(TppCtx c, int b, double* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c.trace_cap * NTRACE) return;
  out[i] = c.trace[(size_t)i * c.Bp + b];
}
__global__ void tpp_cost_sum_kernel(int B, int T, const double* node_cost, double* costs) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double s = 0.;
  for (int t = 0; t <= T; ++t) s += node_cost[(size_t)b * (T + 1) + t];  // shooting.hxx:155-159
  costs[b] = s;
}

template <class M>
void TppBackend<M>::export_scalars(double* d_out) {
  tpp_scalars_kernel<<<nblocks(), 64, 0, stream>>>(c, d_out);
  ++last_launches;
}
template <class M>
void TppBackend<M>::set_reg(double xreg, double ureg, bool set_x, bool set_u) {
  tpp_setreg_kernel<<<nblocks(), 64, 0, stream>>>(c, xreg, ureg, set_x, set_u);
}
template <class M>
void TppBackend<M>::export_trace(int b, double* d_out) {
  if (!c.trace) throw InvalidArg("trace recording is not enabled (cgx_solver_enable_trace)");
  tpp_trace_kernel<<<(c.trace_cap * NTRACE + 63) / 64, 64, 0, stream>>>(c, b, d_out);
}
// trace [cap][NTRACE][Bp] -> [B][nrec][NTRACE]
__global__ void tpp_trace_all_kernel(TppCtx c, int nrec, double* out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (b >= c.B) return;
  for (int j = 0; j < NTRACE; ++j)
    out[((size_t)b * nrec + r) * NTRACE + j] = c.trace[((size_t)r * NTRACE + j) * c.Bp + b];
}
template <class M>
void TppBackend<M>::export_trace_all(int nrec, double* d_out) {
  if (!c.trace) throw InvalidArg("trace recording is not enabled (cgx_solver_enable_trace)");
  if (nrec <= 0) return;
  tpp_trace_all_kernel<<<dim3(nblocks(), nrec), 64, 0, stream>>>(c, nrec, d_out);
  ++last_launches;
}
template <class M>
void TppBackend<M>::problem_calc(const double* d_xs, const double* d_us, int diff, double* d_costs) {
  const size_t nodes = (size_t)c.B * (c.T + 1);
  const int sizes[10] = {1, N, N * N, N * U, N, U, N * N, N * U, U * U, M::NR};
  for (int i = 0; i < 10; ++i)
    if (!node_data[i]) node_data[i] = alloc(nodes * sizes[i]);
  tpp_node_data_kernel<M><<<dim3(nblocks(), c.T + 1), 64, 0, stream>>>(
      c, d_xs, d_us, diff, node_data[0], node_data[1], node_data[2], node_data[3], node_data[4], node_data[5],
      node_data[6], node_data[7], node_data[8], node_data[9]);
  ++last_launches;
  tpp_cost_sum_kernel<<<nblocks(), 64, 0, stream>>>(c.B, c.T, node_data[0], d_costs);
  ++last_launches;
}

}  // namespace

Backend* make_tpp_backend(const ProblemDesc* p, int solver_kind) {
  if (p->kind == 1) {
    if (p->nx != 3 || p->nu != 2) throw InvalidArg("Invalid argument: ActionModelUnicycle has nx = 3, nu = 2");
    if (p->per_problem_params) throw Unsupported("unicycle: per-problem parameters are not supported");
    if (p->nparams != 3) throw InvalidArg("Invalid argument: unicycle params = [dt, w_x, w_u]");
    return new TppBackend<UnicycleModel>(p, solver_kind);
  }
  throw Unsupported("thread-per-problem backend: unsupported model kind");
}

}  // namespace cgx
