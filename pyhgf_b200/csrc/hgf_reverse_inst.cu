// The generated reverse sweep of one topology (pyhgf_b200/revgen.py -> generated/reverse_NAME.h), float32 and float64:
// compile with -DHGF_REV_HEADER=<generated/reverse_NAME.h> -DHGF_REV_TYPE=Rev_NAME -DHGF_REV_FN=reverse_NAME.
#include "hgf_reverse.cuh"
#include HGF_REV_HEADER

namespace hgf {
cudaError_t HGF_REV_FN(const Program& prog, const CallLists& lists, const ScanArgs& args, int dtype, cudaStream_t stream, int* n_launched) {
  ++*n_launched;
  if (dtype == 32) return launch_reverse<float, HGF_REV_TYPE<float>>(prog, lists, args, stream);
  return launch_reverse<double, HGF_REV_TYPE<double>>(prog, lists, args, stream);
}
}  // namespace hgf
