// crocoddyl_b200 — CTA-per-problem backend instantiated for the 4-joint revolute-chain arm (arm.cuh).
#include "cta_backend.cuh"
#include "arm.cuh"

namespace cgx {
Backend* make_cta_arm4_backend(const ProblemDesc* p, int solver_kind) { return new CtaBackend<ArmModel<4> >(p, solver_kind); }
}  // namespace cgx
