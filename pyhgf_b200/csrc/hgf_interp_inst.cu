// One interpreter kernel per translation unit: compile with
// -DHGF_BASE=float|double -DHGF_NTAN=<number of tangent directions, 0 = plain filter>.
#include "hgf_launch.cuh"

namespace hgf {
using Scalar = Dual<HGF_BASE, HGF_NTAN>;  // 0 tangents: the plain filter
template cudaError_t launch_one<Scalar, RuntimeTopo, INTERP_SLOTS, 0>(const Program&, const CallLists&, const ScanArgs&, cudaStream_t);
}
