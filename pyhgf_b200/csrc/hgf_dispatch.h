// Kernel variants: the opcode interpreter (any topology) and the generated
// register-resident specialisations (selected when every topology query of the
// lowered Program has the answer the specialisation was generated with).
#pragma once
#include <cuda_runtime.h>

#include "hgf_program.h"

namespace hgf {

using RunFn = cudaError_t (*)(const Program&, const CallLists&, const ScanArgs&, int dtype, bool want_grad, cudaStream_t,
                              int* n_launched);

// The generated reverse sweep of a topology (hgf_reverse.cuh), when there is one: logp + gradient of a plain call
// (per-instance sums, optional group reduction; no trajectories / pointwise output / fused cross-GPU reduction; every
// response function).
using ReverseFn = cudaError_t (*)(const Program&, const CallLists&, const ScanArgs&, int dtype, cudaStream_t, int* n_launched);

struct KernelVariant {
  const char* name;
  bool (*matches)(const Program&);  // nullptr: accepts every program
  RunFn run;
  ReverseFn reverse = nullptr;
  // the register-resident kernels this variant was built with, per scalar type (variant_bit in hgf_launch_decl.h; 0 for the
  // interpreter, which runs everything): a gradient call whose variant has no gradient kernel of its own would fall to the
  // interpreter's forward mode, so the launcher prefers the generated reverse sweep there whatever the parameter count
  unsigned mask32 = 0, mask64 = 0;
};

// hgf_interp.cu
extern const KernelVariant kInterpreterVariant;
// hgf_static.cu (generated tables): null-terminated list
const KernelVariant* const* static_variants();

const KernelVariant* select_variant(const Program& prog);
// on-demand specialisations (hgf_register_specialisation): consulted before the built-in table
void register_variant(const KernelVariant* v);

// hgf_select.cu: the library's own stream-ordered memory pool on the current device (the reverse sweep's checkpoints)
cudaError_t scratch_pool(cudaMemPool_t* out);

inline cudaError_t run_variant(const KernelVariant& v, const Program& prog, const CallLists& lists, const ScanArgs& args, int dtype,
                               bool want_grad, cudaStream_t stream, int* n_launched) {
  return v.run(prog, lists, args, dtype, want_grad, stream, n_launched);
}

}  // namespace hgf
