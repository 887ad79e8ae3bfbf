// crocoddyl_b200 — CTA-per-problem backend: LQR instantiation and the model factory.
#include "cta_backend.cuh"

namespace cgx {

Backend* make_cta_arm2_backend(const ProblemDesc* p, int solver_kind);
Backend* make_cta_arm4_backend(const ProblemDesc* p, int solver_kind);
Backend* make_cta_arm7_backend(const ProblemDesc* p, int solver_kind);

Backend* make_cta_backend(const ProblemDesc* p, int solver_kind) {
  if (p->kind == 0) return new CtaBackend<LqrModel>(p, solver_kind);
  if (p->kind == 2) {
    switch (p->nu) {
      case 2: return make_cta_arm2_backend(p, solver_kind);
      case 4: return make_cta_arm4_backend(p, solver_kind);
      case 7: return make_cta_arm7_backend(p, solver_kind);
      default:
        throw Unsupported("arm model: kernels are instantiated for nv = 2, 4 and 7 joints (got nv = " +
                          std::to_string(p->nu) + ")");
    }
  }
  throw Unsupported("CTA-per-problem backend: unknown model kind");
}

}  // namespace cgx
