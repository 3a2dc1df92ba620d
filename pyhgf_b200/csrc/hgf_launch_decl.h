// Declarations shared by the dispatching translation units (which never see
// kernel bodies) and the instantiating ones.
#pragma once
#include <cuda_runtime.h>

#include "hgf_math.cuh"
#include "hgf_program.h"

namespace hgf {

struct RuntimeTopo;

template <class T, class Topo, int NSLOTS>
cudaError_t launch_one(const Program& prog, const CallLists& lists, const ScanArgs& args, cudaStream_t stream);

// Forward filter: one launch.  Gradient: forward-mode passes of up to four
// tangent directions each; the first pass also produces the non-gradient
// outputs.  NARROW = false restricts the passes to width 4 (fewer kernels to
// build; used for float64).
template <class B, class Topo, int NSLOTS, bool NARROW>
inline cudaError_t launch_scan(const Program& prog, const CallLists& lists, const ScanArgs& args, bool want_grad,
                               cudaStream_t stream, int* n_launched) {
  if (!want_grad) {
    ++*n_launched;
    return launch_one<Dual<B, 0>, Topo, NSLOTS>(prog, lists, args, stream);
  }
  ScanArgs a = args;
  int done = 0;
  while (done < lists.n_params) {
    const int rem = lists.n_params - done;
    a.tangent_offset = done;
    cudaError_t err;
    if (rem >= 4 || !NARROW) { err = launch_one<Dual<B, 4>, Topo, NSLOTS>(prog, lists, a, stream); done += 4; }
    else if (rem == 3) { err = launch_one<Dual<B, NARROW ? 3 : 4>, Topo, NSLOTS>(prog, lists, a, stream); done += 3; }
    else if (rem == 2) { err = launch_one<Dual<B, NARROW ? 2 : 4>, Topo, NSLOTS>(prog, lists, a, stream); done += 2; }
    else { err = launch_one<Dual<B, NARROW ? 1 : 4>, Topo, NSLOTS>(prog, lists, a, stream); done += 1; }
    ++*n_launched;
    if (err != cudaSuccess) return err;
    // later passes only add gradient columns
    a.trajectories = nullptr; a.last_attributes = nullptr; a.surprise_sum = nullptr; a.surprise_pointwise = nullptr; a.reduced = nullptr;
  }
  return cudaSuccess;
}

}  // namespace hgf
