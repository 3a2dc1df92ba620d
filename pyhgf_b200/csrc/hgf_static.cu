// Table of the generated register-resident specialisations (pyhgf_b200/codegen.py
// writes generated/topo_*.h and generated/static_table.inc).  Kernel bodies are
// instantiated in hgf_static_inst.cu, one (topology, scalar type) per translation unit.
#include "hgf_static_run.h"

#if __has_include(<generated/static_includes.inc>)
#include <generated/static_includes.inc>
#endif

namespace hgf {

// reverse sweeps (generated/reverse_table.inc lists the topologies that have one; instantiated in hgf_reverse_inst.cu)
#define HGF_STATIC_REVERSE(name) \
  cudaError_t reverse_##name(const Program&, const CallLists&, const ScanArgs&, int, cudaStream_t, int*);
#if __has_include(<generated/reverse_table.inc>)
#include <generated/reverse_table.inc>
#endif
#undef HGF_STATIC_REVERSE
template <int> struct no_reverse { static constexpr ReverseFn fn = nullptr; };
#define HGF_STATIC_REVERSE(name) static constexpr ReverseFn kReverse_##name = reverse_##name;
#if __has_include(<generated/reverse_table.inc>)
#include <generated/reverse_table.inc>
#endif
#undef HGF_STATIC_REVERSE

#define HGF_STATIC_TOPO(name, mask32, mask64, rev)                                                   \
  static const KernelVariant kVariant_##name = {"static:" #name, topo_matches<Topo_##name>,          \
                                                run_static_any<Topo_##name, mask32, mask64>, rev, mask32, mask64};
#if __has_include(<generated/static_table.inc>)
#include <generated/static_table.inc>
#endif
#undef HGF_STATIC_TOPO

static const KernelVariant* const kStatic[] = {
#define HGF_STATIC_TOPO(name, mask32, mask64, rev) &kVariant_##name,
#if __has_include(<generated/static_table.inc>)
#include <generated/static_table.inc>
#endif
#undef HGF_STATIC_TOPO
    nullptr};

const KernelVariant* const* static_variants() { return kStatic; }

}  // namespace hgf
