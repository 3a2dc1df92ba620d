// One interpreter kernel per translation unit: compile with
// -DHGF_BASE=float|double -DHGF_NTAN=<number of tangent directions, 0 = plain filter>.
#include <type_traits>

#include "hgf_launch.cuh"

namespace hgf {
using Scalar = std::conditional_t<HGF_NTAN == 0, HGF_BASE, Dual<HGF_BASE, (HGF_NTAN > 0 ? HGF_NTAN : 1)>>;
template cudaError_t launch_one<Scalar, RuntimeTopo, INTERP_SLOTS>(const Program&, const CallLists&, const ScanArgs&, cudaStream_t);
}
