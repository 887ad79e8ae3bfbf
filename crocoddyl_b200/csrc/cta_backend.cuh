// crocoddyl_b200 — host driver of the CTA-per-problem kernels (cta.cuh), a template over the node model. It is
// instantiated in one translation unit per model (cta_backend.cu: LQR and the factory; cta_arm{2,4,7}.cu: the
// revolute-chain arm) so that the models compile in parallel.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <vector>
#include "backend.hpp"
#include "cta.cuh"

namespace cgx {

namespace {

template <class M>
struct CtaBackend : Backend {
  CtaCtx c;
  std::vector<void*> allocs;
  bool solver_allocated = false, data_allocated = false;
  int threads = 128;
  size_t smem = 0;
  bool phase_split = M::kPhaseSplit;  // CGX_CTA_FUSED=1 / CGX_CTA_PHASES=1 override (development A/B knob)
  TrialSlots slots;       // spare trial buffers of the concurrent line search (allocated on first use)
  int wide_mode = -1;     // -1 automatic, 0 never, 1 always (CGX_CTA_WIDE, development / test knob)
  int num_sms = 148;
  int* d_active = nullptr;
  int* h_active = nullptr;
  cudaEvent_t ring_ev[4] = {nullptr, nullptr, nullptr, nullptr};

  explicit CtaBackend(const ProblemDesc* p_, int sk) {
    p = p_;
    solver_kind = sk;
    std::memset(&c, 0, sizeof(c));
    c.B = p->B;
    c.Bp = p->B;
    c.T = p->T;
    c.n = p->nx;
    c.m = p->nu;
    c.fddp = sk == 1 || sk == 3;  // SolverFDDP / SolverBoxFDDP
    c.box = sk >= 2;              // SolverBoxDDP / SolverBoxFDDP
    c.model_kind = p->kind;
    c.params = p->d_params;
    c.tparams = p->d_tparams;
    c.u_lb = p->d_ulb;
    c.u_ub = p->d_uub;
    c.has_limits = p->d_haslim;
    c.nparams = p->nparams;
    c.per_problem = p->per_problem_params;
    c.x0 = p->d_x0;
    CGX_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    for (int i = 0; i < 4; ++i) CGX_CUDA(cudaEventCreate(&ev[i]));
    const int tn = (c.n + 3) / 4;
    threads = ((tn * tn + 31) / 32) * 32;
    if (threads < M::kMinThreads) threads = M::kMinThreads;
    if (threads > M::kMaxThreads) threads = M::kMaxThreads;
    // dense nodes on the FP64 tensor-core path: one warp per 24 x 24 tile of the widest phase ([Fx Fu]^T Vxx)
    if (c.n >= 48 && !cta_is_small(c.n, c.m)) {
      threads = 32 * cta_mma_tiles(c.n + c.m, c.n);
      if (threads < M::kMinThreads) threads = M::kMinThreads;
      if (threads > M::kMaxThreads) threads = M::kMaxThreads;
    }
    // small nodes: the backward pass is one warp's work (cta_small.cuh), a second warp only halves the resident
    // problems per SM (measured, LQR n = 12, B = 4096: see profiles/r01_riccati_sweep_v3.json)
    if (cta_is_small(c.n, c.m)) threads = 32;
    smem = cta_smem_bytes(c.n, c.m);
    int dev = 0, maxsm = 0;
    CGX_CUDA(cudaGetDevice(&dev));
    CGX_CUDA(cudaDeviceGetAttribute(&maxsm, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    if (smem > (size_t)maxsm)
      throw Unsupported("problem too large for one thread block's shared memory (nx = " + std::to_string(c.n) + ")");
    CGX_CUDA(cudaFuncSetAttribute(cta_solve_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CGX_CUDA(cudaFuncSetAttribute(cta_phase_direction_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CGX_CUDA(cudaFuncSetAttribute(cta_phase_forward_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CGX_CUDA(cudaFuncSetAttribute(cta_phase_forward_wide_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    std::memset(&slots, 0, sizeof(slots));
    CGX_CUDA(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
    if (const char* e = getenv("CGX_CTA_WIDE")) wide_mode = atoi(e);
    if (const char* e = getenv("CGX_CTA_FUSED")) if (e[0] == '1') phase_split = false;
    if (const char* e = getenv("CGX_CTA_PHASES")) if (e[0] == '1') phase_split = true;
    CGX_CUDA(cudaFuncSetAttribute(cta_direction_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CGX_CUDA(cudaFuncSetAttribute(cta_trystep_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CGX_CUDA(cudaFuncSetAttribute(cta_rollout_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  ~CtaBackend() override {
    for (void* q : allocs) cudaFree(q);
    if (h_active) cudaFreeHost(h_active);
    for (int i = 0; i < 4; ++i)
      if (ring_ev[i]) cudaEventDestroy(ring_ev[i]);
    for (int i = 0; i < 4; ++i)
      if (ev[i]) cudaEventDestroy(ev[i]);
    if (stream) cudaStreamDestroy(stream);
  }
  double* alloc(size_t n) {
    double* q = dalloc<double>(n);
    CGX_CUDA(cudaMemsetAsync(q, 0, (n ? n : 1) * sizeof(double), stream));
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
  int sblocks() const { return (c.B + 127) / 128; }
  void x0_changed() override {}

  void ensure_data() {
    if (data_allocated) return;
    const size_t B = c.B, T = c.T, n = c.n, m = c.m;
    c.d_cost = alloc(B * (T + 1));
    c.d_xnext = alloc(B * (T + 1) * n);
    c.d_Lx = alloc(B * (T + 1) * n);
    c.d_Lu = alloc(B * (T + 1) * (m ? m : 1));
    if (M::kPerNodeMats) {
      c.d_Fx = alloc(B * (T + 1) * n * n);
      c.d_Fu = alloc(B * (T + 1) * n * m);
      c.d_Lxx = alloc(B * (T + 1) * n * n);
      c.d_Lxu = alloc(n * m);  // identically zero for the arm's cost set: one shared block
      c.d_Luu = alloc(B * (T + 1) * m * m);
    }
    data_allocated = true;
  }

  void configure() override {
    c.o = o;
    const size_t B = c.B, T = c.T, n = c.n, m = c.m;
    ensure_data();
    if (!solver_allocated) {
      for (int i = 0; i < 2; ++i) {
        c.xs[i] = alloc(B * (T + 1) * n);
        c.us[i] = alloc(B * T * m);
      }
      c.fs = alloc(B * (T + 1) * n);
      c.K = alloc(B * T * m * n);
      c.k = alloc(B * T * m);
      c.Vx = alloc(B * (T + 1) * n);
      c.Vxx = alloc(B * (T + 1) * n * n);
      c.w = alloc(B * (T + 1) * n);
      c.scal = alloc((size_t)S_NSLOTS * c.Bp);
      int* is = dalloc<int>((size_t)I_NSLOTS * c.Bp);
      allocs.push_back(is);
      c.iscal = is;
      CGX_CUDA(cudaMemsetAsync(is, 0, (size_t)I_NSLOTS * c.Bp * sizeof(int), stream));
      std::vector<double> h((size_t)S_NSLOTS * c.Bp, 0.);
      for (size_t b = 0; b < (size_t)c.Bp; ++b) {
        h[(size_t)S_XREG * c.Bp + b] = NAN;  // solver-base.cpp:14-25
        h[(size_t)S_UREG * c.Bp + b] = NAN;
        h[(size_t)S_STEP * c.Bp + b] = 1.;
      }
      CGX_CUDA(cudaMemcpyAsync(c.scal, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice, stream));
      CGX_CUDA(cudaStreamSynchronize(stream));
      solver_allocated = true;
    }
    if (c.box && !c.Quu_inv) {
      c.Quu_inv = alloc(B * T * m * m);
      c.box_work = alloc(B * (size_t)box_gains_work_doubles((int)m));
      int* iw = dalloc<int>(B * (size_t)box_gains_work_ints((int)m) + 1);
      allocs.push_back(iw);
      c.box_iwork = iw;
    }
    if (full && !c.Qxx) {
      c.Qxx = alloc(B * T * n * n);
      c.Qxu = alloc(B * T * n * m);
      c.Quu = alloc(B * T * m * m);
      c.Qx = alloc(B * T * n);
      c.Qu = alloc(B * T * m);
    }
    c.full = full;
    if (trace_on && (!c.trace || c.trace_cap != trace_cap)) {
      release(c.trace);
      c.trace = alloc(B * (size_t)trace_cap * NTRACE);
      c.trace_cap = trace_cap;
    }
    if (!trace_on) {
      release(c.trace);
      c.trace_cap = 0;
    }
  }

  void set_candidate(const double* d_xs, const double* d_us, int is_feasible) override {
    const size_t B = c.B, T = c.T, n = c.n, m = c.m;
    cta_reset_kernel<<<sblocks(), 128, 0, stream>>>(c, is_feasible);
    ++last_launches;
    if (d_xs)
      CGX_CUDA(cudaMemcpyAsync(c.xs[0], d_xs, B * (T + 1) * n * sizeof(double), cudaMemcpyDeviceToDevice, stream));
    else
      CGX_CUDA(cudaMemsetAsync(c.xs[0], 0, B * (T + 1) * n * sizeof(double), stream));
    if (d_us)
      CGX_CUDA(cudaMemcpyAsync(c.us[0], d_us, B * T * m * sizeof(double), cudaMemcpyDeviceToDevice, stream));
    else
      CGX_CUDA(cudaMemsetAsync(c.us[0], 0, B * T * m * sizeof(double), stream));
  }

  void solve(int maxiter, int is_feasible, double reginit) override {
    c.o = o;
    CGX_CUDA(cudaEventRecord(ev[1], stream));
    if (phase_split) {
      solve_phases(maxiter, is_feasible, reginit);
    } else {
      cta_solve_kernel<M><<<c.B, threads, smem, stream>>>(c, maxiter, is_feasible, reginit);
      ++last_launches;
    }
    CGX_CUDA(cudaEventRecord(ev[2], stream));
    CGX_CUDA(cudaGetLastError());
  }
  void ensure_trial_slots() {
    if (slots.res) return;
    const size_t B = c.B, T = c.T, n = c.n, m = c.m;
    for (int g = 0; g < CTA_LS_WIDTH - 1; ++g) {
      slots.spare[g].xs = alloc(B * (T + 1) * n);
      slots.spare[g].us = alloc(B * T * m);
      slots.spare[g].xnext = alloc(B * (T + 1) * n);
      slots.spare[g].cost = alloc(B * (T + 1));
    }
    slots.res = alloc((size_t)CTA_LS_WIDTH * 3 * c.Bp);
  }
  // Phase-split solve (cta.cuh): alternate the direction and the line-search launches over the whole batch until every
  // problem has returned. The count of unfinished problems a direction launch saw comes back through pinned memory
  // two rounds late, so the host never stalls the stream; the two surplus rounds are empty launches.
  void solve_phases(int maxiter, int is_feasible, double reginit) {
    constexpr int RING = 4;
    if (!d_active) {
      d_active = dalloc<int>(RING);
      allocs.push_back(d_active);
      CGX_CUDA(cudaMallocHost((void**)&h_active, RING * sizeof(int)));
      for (int i = 0; i < RING; ++i) CGX_CUDA(cudaEventCreateWithFlags(&ring_ev[i], cudaEventDisableTiming));
    }
    cta_phase_init_kernel<<<sblocks(), 128, 0, stream>>>(c, maxiter, is_feasible, reginit);
    ++last_launches;
    // upper bound of rounds: every round an unfinished problem either runs a backward pass (at most one per step of
    // the regularisation from regmin to regmax, ddp.cpp:312-318) or a trial of the line search
    const double span = std::log(std::max(c.o.regmax, 1e-300) / std::max(c.o.regmin, 1e-300)) / std::log(c.o.regfactor);
    const long max_retries = (long)std::ceil(std::max(span, 0.)) + 2;
    const long max_rounds = (long)maxiter * (c.o.nalphas + max_retries) + RING;
    // rollout by one thread per problem (cta_forward_thread) while many problems are unfinished: a warp instruction then
    // advances 32 rollouts instead of one. A lone rollout is ~1.5x slower that way, so the warp version takes over for
    // the tail (measured, arm C4: 143 ms with warps, 187 ms with threads). The two node functions differ at rounding
    // level (arm_warp.cuh); the sequence of counts, and so the schedule, is deterministic for a given batch.
    int tpp_min = 1536;
    if (const char* e = getenv("CGX_CTA_TPP_FORWARD_MIN")) tpp_min = atoi(e);
    long unfinished = c.B;  // last count that came back (never smaller than the truth)
    long last_round = -1;
    bool finished = false;
    // development aid: CGX_CTA_PHASE_TIMING=1 prints, per round, the unfinished count the host knew and the duration
    // of the three phases (timing events on the solve stream; nothing is recorded otherwise)
    static const bool phase_timing = getenv("CGX_CTA_PHASE_TIMING") && getenv("CGX_CTA_PHASE_TIMING")[0] == '1';
    std::vector<cudaEvent_t> pt_ev;
    std::vector<long> pt_unf;
    auto pt_mark = [&]() {
      if (!phase_timing) return;
      cudaEvent_t e;
      CGX_CUDA(cudaEventCreate(&e));
      CGX_CUDA(cudaEventRecord(e, stream));
      pt_ev.push_back(e);
    };
    for (long r = 0; r < max_rounds; ++r) {
      const int slot = (int)(r % RING);
      if (r >= RING - 1) {  // result of round r - (RING - 1)
        const int old = (int)((r - (RING - 1)) % RING);
        CGX_CUDA(cudaEventSynchronize(ring_ev[old]));
        if (h_active[old] == 0) {
          finished = true;
          break;
        }
        unfinished = h_active[old];
      }
      // concurrent line search while all trials of the unfinished problems fit the GPU in one wave (cta.cuh)
      const bool wide = wide_mode == 1 || (wide_mode < 0 && unfinished * CTA_LS_WIDTH <= (long)num_sms * M::kFwdMinBlocks);
      if (wide) ensure_trial_slots();
      const bool tpp_forward = M::kN > 0 && unfinished >= tpp_min;
      CGX_CUDA(cudaMemsetAsync(d_active + slot, 0, sizeof(int), stream));
      pt_unf.push_back(unfinished);
      pt_mark();
      if constexpr (M::kNodeKernel) {
        const long nodes = (long)c.B * (c.T + 1);
        cta_phase_nodes_kernel<M><<<(unsigned)((nodes + 63) / 64), 64, 0, stream>>>(c);
        ++last_launches;
      }
      pt_mark();
      cta_phase_direction_kernel<M><<<c.B, threads, smem, stream>>>(c, d_active + slot);
      pt_mark();
      if (wide) {
        if (tpp_forward)
          cta_phase_forward_wide_tpp_kernel<M><<<dim3((c.B + 31) / 32, CTA_LS_WIDTH), 32, 0, stream>>>(c, slots);
        else
          cta_phase_forward_wide_kernel<M><<<dim3(c.B, CTA_LS_WIDTH), threads, cta_blockwide_smem_bytes(c.n, c.m), stream>>>(c, slots);
        cta_phase_select_kernel<<<c.B, 128, 0, stream>>>(c, slots, maxiter);
        ++last_launches;
      } else if (tpp_forward) {
        cta_phase_forward_tpp_kernel<M><<<(c.B + 31) / 32, 32, 0, stream>>>(c, maxiter);
      } else {
        cta_phase_forward_kernel<M><<<c.B, threads, cta_blockwide_smem_bytes(c.n, c.m), stream>>>(c, maxiter);
      }
      last_launches += 2;
      pt_mark();
      CGX_CUDA(cudaMemcpyAsync(h_active + slot, d_active + slot, sizeof(int), cudaMemcpyDeviceToHost, stream));
      CGX_CUDA(cudaEventRecord(ring_ev[slot], stream));
      last_round = r;
    }
    if (phase_timing) {
      CGX_CUDA(cudaStreamSynchronize(stream));
      double tot[3] = {0., 0., 0.};
      for (size_t r = 0; r * 4 + 3 < pt_ev.size(); ++r) {
        float ms[3];
        for (int k = 0; k < 3; ++k) {
          CGX_CUDA(cudaEventElapsedTime(&ms[k], pt_ev[r * 4 + k], pt_ev[r * 4 + k + 1]));
          tot[k] += ms[k];
        }
        fprintf(stderr, "[cgx phases] round %3zu unfinished<=%6ld nodes %.3f direction %.3f forward %.3f ms\n", r, pt_unf[r],
                ms[0], ms[1], ms[2]);
      }
      fprintf(stderr, "[cgx phases] total nodes %.2f direction %.2f forward %.2f ms over %zu rounds\n", tot[0], tot[1], tot[2],
              pt_ev.size() / 4);
      for (cudaEvent_t e : pt_ev) cudaEventDestroy(e);
    }
    if (!finished) {  // the round bound was hit: make sure that every problem has returned (backstop)
      CGX_CUDA(cudaStreamSynchronize(stream));
      if (last_round >= 0 && h_active[last_round % RING] != 0)
        throw CudaError("phase-split solve: " + std::to_string(h_active[last_round % RING]) +
                        " problems had not returned when the round bound was reached");
    }
  }
  void compute_direction(int recalc) override {
    c.o = o;
    CGX_CUDA(cudaEventRecord(ev[1], stream));
    cta_direction_kernel<M><<<c.B, threads, smem, stream>>>(c, 2, recalc);
    CGX_CUDA(cudaEventRecord(ev[2], stream));
    ++last_launches;
    CGX_CUDA(cudaGetLastError());
  }
  void try_step(double alpha) override {
    c.o = o;
    CGX_CUDA(cudaEventRecord(ev[1], stream));
    cta_trystep_kernel<M><<<c.B, threads, smem, stream>>>(c, alpha);
    CGX_CUDA(cudaEventRecord(ev[2], stream));
    ++last_launches;
    CGX_CUDA(cudaGetLastError());
  }

  struct Fld {
    double* a;
    double* b;
    int flip, size, nt;
    bool dbl;
  };
  Fld fld(const std::string& nme) {
    const int T = c.T, n = c.n, m = c.m;
    if (nme == "xs") return {c.xs[0], c.xs[1], 0, n, T + 1, true};
    if (nme == "xs_try") return {c.xs[0], c.xs[1], 1, n, T + 1, true};
    if (nme == "us") return {c.us[0], c.us[1], 0, m, T, true};
    if (nme == "us_try") return {c.us[0], c.us[1], 1, m, T, true};
    if (nme == "fs") return {c.fs, c.fs, 0, n, T + 1, false};
    if (nme == "K") return {c.K, c.K, 0, m * n, T, false};
    if (nme == "k") return {c.k, c.k, 0, m, T, false};
    if (nme == "Vx") return {c.Vx, c.Vx, 0, n, T + 1, false};
    if (nme == "Vxx") return {c.Vxx, c.Vxx, 0, n * n, T + 1, false};
    if (nme == "Quu_inv") {  // SolverBoxDDP::get_Quu_inv (box-ddp.hpp)
      if (!c.Quu_inv) throw Unsupported("Quu_inv is kept by the box solvers only");
      return {c.Quu_inv, c.Quu_inv, 0, m * m, T, false};
    }
    if (nme == "Qxx" || nme == "Qxu" || nme == "Quu" || nme == "Qx" || nme == "Qu") {
      if (!c.full || !c.Qxx)
        throw Unsupported("field " + nme + " is not kept with cgx_solver_set_storage(full = 0)");
      if (nme == "Qxx") return {c.Qxx, c.Qxx, 0, n * n, T, false};
      if (nme == "Qxu") return {c.Qxu, c.Qxu, 0, n * m, T, false};
      if (nme == "Quu") return {c.Quu, c.Quu, 0, m * m, T, false};
      if (nme == "Qx") return {c.Qx, c.Qx, 0, n, T, false};
      return {c.Qu, c.Qu, 0, m, T, false};
    }
    throw NotFound("unknown field " + nme);
  }
  FieldInfo field_info(const std::string& name) override {
    Fld f = fld(name);
    return {f.size, f.nt};
  }
  void copy_sel(double* d_user, const Fld& f, int b0, int nb, int to_field) {
    const size_t per = (size_t)f.size * f.nt;
    if (per == 0) return;
    int gx = (int)((per + 255) / 256);
    if (gx > 64) gx = 64;
    cta_copy_sel_kernel<<<dim3(gx, nb), 256, 0, stream>>>(d_user, f.a, f.b, f.dbl ? c.iscal + (size_t)I_CUR * c.Bp : nullptr,
                                                          f.flip, per, b0, to_field);
    ++last_launches;
  }
  void export_field(const std::string& name, int b0, int nb, double* d_out) override { copy_sel(d_out, fld(name), b0, nb, 0); }
  void import_field(const std::string& name, const double* d_in) override {
    if (name != "xs" && name != "us") throw Unsupported("only xs and us can be set");
    copy_sel(const_cast<double*>(d_in), fld(name), 0, c.B, 1);
  }
  void export_scalars(double* d_out) override {
    cta_scalars_kernel<<<sblocks(), 128, 0, stream>>>(c, d_out);
    ++last_launches;
  }
  void export_status(int which, int* d_out) override {
    CGX_CUDA(cudaMemcpyAsync(d_out, c.iscal + (size_t)(which ? I_FSTATUS : I_BSTATUS) * c.Bp, c.B * sizeof(int),
                             cudaMemcpyDeviceToDevice, stream));
  }
  void set_reg(double xreg, double ureg, bool set_x, bool set_u) override {
    cta_setreg_kernel<<<sblocks(), 128, 0, stream>>>(c, xreg, ureg, set_x, set_u);
  }
  void export_trace(int b, double* d_out) override {
    if (!c.trace) throw InvalidArg("trace recording is not enabled (cgx_solver_enable_trace)");
    CGX_CUDA(cudaMemcpyAsync(d_out, c.trace + (size_t)b * c.trace_cap * NTRACE, (size_t)c.trace_cap * NTRACE * sizeof(double),
                             cudaMemcpyDeviceToDevice, stream));
  }

  void export_trace_all(int nrec, double* d_out) override {
    if (!c.trace) throw InvalidArg("trace recording is not enabled (cgx_solver_enable_trace)");
    if (nrec <= 0) return;
    CGX_CUDA(cudaMemcpy2DAsync(d_out, (size_t)nrec * NTRACE * sizeof(double), c.trace, (size_t)c.trace_cap * NTRACE * sizeof(double),
                               (size_t)nrec * NTRACE * sizeof(double), c.B, cudaMemcpyDeviceToDevice, stream));
  }

  // ---- problem-level API ----
  void problem_calc(const double* d_xs, const double* d_us, int diff, double* d_costs) override {
    ensure_data();
    cta_problem_kernel<M><<<c.B, 128, 0, stream>>>(c, d_xs, d_us, diff, d_costs);
    ++last_launches;
  }
  void problem_rollout(const double* d_us, double* d_xs) override {
    ensure_data();
    cta_rollout_kernel<M><<<c.B, threads, smem, stream>>>(c, d_us, d_xs);
    ++last_launches;
  }
  void problem_quasi_static(const double* d_x, double* d_u) override { M::quasi_static(*this, c, d_x, d_u, stream); }
  FieldInfo data_info(const std::string& name) override {
    const int T = c.T, n = c.n, m = c.m;
    if (name == "cost") return {1, T + 1};
    if (name == "xnext") return {n, T + 1};
    if (name == "Fx") return {n * n, T + 1};
    if (name == "Fu") return {n * m, T + 1};
    if (name == "Lx") return {n, T + 1};
    if (name == "Lu") return {m, T + 1};
    if (name == "Lxx") return {n * n, T + 1};
    if (name == "Lxu") return {n * m, T + 1};
    if (name == "Luu") return {m * m, T + 1};
    throw NotFound("unknown data field " + name);
  }
  void export_data(const std::string& name, int b, int t, double* d_out) override {
    ensure_data();
    FieldInfo fi = data_info(name);
    const size_t node = (size_t)b * (c.T + 1) + t;
    const double* src = nullptr;
    if (name == "cost") src = c.d_cost + node;
    else if (name == "xnext") src = c.d_xnext + node * c.n;
    else if (name == "Lx") src = c.d_Lx + node * c.n;
    else if (name == "Lu") src = c.d_Lu + node * c.m;
    else src = M::node_matrix(c, p, name, b, t);
    CGX_CUDA(cudaMemcpyAsync(d_out, src, fi.size * sizeof(double), cudaMemcpyDeviceToDevice, stream));
  }
};

}  // namespace

}  // namespace cgx
