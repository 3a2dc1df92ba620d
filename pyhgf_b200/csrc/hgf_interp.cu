// Opcode interpreter: runs any lowered Program.  The instance's attribute
// vector is a per-thread local array (dynamically indexed -> local memory,
// which the hardware interleaves per thread, i.e. [slot][thread] in L1).
// Kernel bodies are instantiated in hgf_interp_inst.cu, one per scalar type.
#include "hgf_dispatch.h"
#include "hgf_launch_decl.h"

namespace hgf {

static cudaError_t run_interpreter(const Program& prog, const CallLists& lists, const ScanArgs& args, int dtype, bool want_grad,
                                   cudaStream_t stream, int* n_launched) {
  if (prog.n_slots > INTERP_SLOTS) return cudaErrorInvalidValue;  // rejected earlier by hgf_plan_create
  // instantiated in hgf_interp_inst.cu: float PF = 0..4, double PF = 0 and 4; no response-only tangents
  if (dtype == 32) return launch_scan<float, RuntimeTopo, INTERP_SLOTS, 0x1fu>(prog, lists, args, want_grad, stream, n_launched);
  return launch_scan<double, RuntimeTopo, INTERP_SLOTS, 0x11u>(prog, lists, args, want_grad, stream, n_launched);
}

const KernelVariant kInterpreterVariant = {"interpreter", nullptr, run_interpreter};

}  // namespace hgf
