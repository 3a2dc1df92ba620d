// One register-resident kernel per translation unit: compile with
//   -DHGF_TOPO_HEADER=<generated/topo_NAME.h> -DHGF_TOPO_TYPE=Topo_NAME
//   -DHGF_BASE=float|double -DHGF_NTAN=<filter tangent directions, 0 = plain filter>
//   -DHGF_NRESP=<response-only tangent directions, 0 or 1>.
#include "hgf_launch.cuh"
#include HGF_TOPO_HEADER

namespace hgf {
using Scalar = Dual<HGF_BASE, HGF_NTAN>;  // 0 tangents: the plain filter
template cudaError_t launch_one<Scalar, HGF_TOPO_TYPE, HGF_TOPO_TYPE::kSlots, HGF_NRESP>(const Program&, const CallLists&, const ScanArgs&,
                                                                             cudaStream_t);
#if HGF_NTAN > 0
// the flavour for calls that differentiate no structural attribute (hgf_ops.cuh::attr, launch_scan picks it)
template cudaError_t launch_one<Scalar, ConstAttributes<HGF_TOPO_TYPE>, HGF_TOPO_TYPE::kSlots, HGF_NRESP>(const Program&, const CallLists&,
                                                                                              const ScanArgs&, cudaStream_t);
#endif
}
