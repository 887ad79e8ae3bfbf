// placeholder until the CTA-per-problem backend lands
#include "backend.hpp"
namespace cgx {
Backend* make_cta_backend(const ProblemDesc*, int) { throw Unsupported("CTA backend not built yet"); }
}
