// TEST INFRASTRUCTURE — CPU oracle (see linalg.hpp header). Not part of the product.
#pragma once
#include "models.hpp"
namespace oracle {
inline void make_arm_models(const double*, long, int, int, std::shared_ptr<ActionModel>&, std::shared_ptr<ActionModel>&) {
  throw std::invalid_argument("arm model not built yet");
}
}
namespace oracle { inline int arm_op(const double*, long, const char*, const double*, double*) { return -1; } }
