#include <cstring>

#include "hgf_dispatch.h"

namespace hgf {

const KernelVariant* select_variant(const Program& prog) {
  for (const KernelVariant* const* v = static_variants(); *v; ++v)
    if ((*v)->matches && (*v)->matches(prog)) return *v;
  return &kInterpreterVariant;
}

}  // namespace hgf
