// Table entry of ONE on-demand specialisation (pyhgf_b200/specialise.py): compile with
//   -DHGF_TOPO_HEADER=<topo_NAME.h> -DHGF_TOPO_TYPE=Topo_NAME -DHGF_TOPO_LABEL="specialised:NAME"
//   -DHGF_MASK32=<instantiated float kernels> -DHGF_MASK64=<instantiated double kernels>
// (variant_bit in hgf_launch_decl.h); the kernels themselves come from hgf_static_inst.cu, one per translation
// unit, exactly as for the built-in topologies.  With -DHGF_REV_FN=reverse_NAME the object also carries the topology's
// generated reverse sweep (hgf_reverse_inst.cu compiled with the same name).  Loaded with hgf_register_specialisation.
#include "../../include/pyhgf_cuda.h"
#include "hgf_static_run.h"
#include HGF_TOPO_HEADER

namespace hgf {
#ifdef HGF_REV_FN
cudaError_t HGF_REV_FN(const Program&, const CallLists&, const ScanArgs&, int, cudaStream_t, int*);
static constexpr ReverseFn kSpecialisedReverse = HGF_REV_FN;
#else
static constexpr ReverseFn kSpecialisedReverse = nullptr;
#endif
static const KernelVariant kSpecialised = {HGF_TOPO_LABEL, topo_matches<HGF_TOPO_TYPE>, run_static_any<HGF_TOPO_TYPE, HGF_MASK32, HGF_MASK64>,
                                           kSpecialisedReverse, HGF_MASK32, HGF_MASK64};
}

extern "C" const void* hgf_specialisation_variant() { return &hgf::kSpecialised; }
extern "C" int hgf_specialisation_abi() { return HGF_ABI_VERSION; }
