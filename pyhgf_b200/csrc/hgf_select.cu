#include <cstdlib>
#include <mutex>
#include <vector>

#include "hgf_dispatch.h"

namespace hgf {

// HGF_FORCE_INTERPRETER=1 (read at plan creation) routes every network through the opcode interpreter:
// used by the tests to run both kernel families on the same inputs.
static std::mutex g_registered_mutex;
static std::vector<const KernelVariant*>& registered() {
  static std::vector<const KernelVariant*> v;
  return v;
}

void register_variant(const KernelVariant* v) {
  std::lock_guard<std::mutex> lock(g_registered_mutex);
  for (const KernelVariant* r : registered())
    if (r == v) return;
  registered().push_back(v);
}

const KernelVariant* select_variant(const Program& prog) {
  const char* force = std::getenv("HGF_FORCE_INTERPRETER");
  if (force && force[0] == '1') return &kInterpreterVariant;
  {
    std::lock_guard<std::mutex> lock(g_registered_mutex);
    for (auto it = registered().rbegin(); it != registered().rend(); ++it)  // newest first
      if ((*it)->matches && (*it)->matches(prog)) return *it;
  }
  for (const KernelVariant* const* v = static_variants(); *v; ++v)
    if ((*v)->matches && (*v)->matches(prog)) return *v;
  return &kInterpreterVariant;
}

}  // namespace hgf
