#include <cstdlib>

#include "hgf_dispatch.h"

namespace hgf {

// HGF_FORCE_INTERPRETER=1 (read at plan creation) routes every network through the opcode interpreter:
// used by the tests to run both kernel families on the same inputs.
const KernelVariant* select_variant(const Program& prog) {
  const char* force = std::getenv("HGF_FORCE_INTERPRETER");
  if (force && force[0] == '1') return &kInterpreterVariant;
  for (const KernelVariant* const* v = static_variants(); *v; ++v)
    if ((*v)->matches && (*v)->matches(prog)) return *v;
  return &kInterpreterVariant;
}

}  // namespace hgf
