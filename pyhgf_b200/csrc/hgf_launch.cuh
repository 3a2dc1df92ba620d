// Definition of launch_one (instantiated explicitly, one kernel family per
// translation unit so that the build parallelises).
#pragma once
#include "hgf_launch_decl.h"
#include "hgf_scan.cuh"

namespace hgf {

template <class T, class Topo, int NSLOTS, int PR>
cudaError_t launch_one(const Program& prog, const CallLists& lists, const ScanArgs& args, cudaStream_t stream) {
  const long long blocks = (args.n_instances + SCAN_THREADS - 1) / SCAN_THREADS;
  if constexpr (Topo::is_static) {
    if (!args.time_steps) {  // dt = 1 folds away (and with it log dt and the per-step load)
      scan_kernel<T, Topo, NSLOTS, PR, true><<<static_cast<unsigned>(blocks), SCAN_THREADS, 0, stream>>>(prog, lists, args);
      return cudaGetLastError();
    }
  }
  scan_kernel<T, Topo, NSLOTS, PR, false><<<static_cast<unsigned>(blocks), SCAN_THREADS, 0, stream>>>(prog, lists, args);
  return cudaGetLastError();
}

}  // namespace hgf
