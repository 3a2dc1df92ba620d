// Definition of launch_one (instantiated explicitly, one kernel per translation
// unit so that the build parallelises) and the pass logic of the gradient.
#pragma once
#include "hgf_launch_decl.h"
#include "hgf_scan.cuh"

namespace hgf {

template <class T, class Topo, int NSLOTS>
cudaError_t launch_one(const Program& prog, const CallLists& lists, const ScanArgs& args, cudaStream_t stream) {
  const long long blocks = (args.n_instances + SCAN_THREADS - 1) / SCAN_THREADS;
  scan_kernel<T, Topo, NSLOTS><<<static_cast<unsigned>(blocks), SCAN_THREADS, 0, stream>>>(prog, lists, args);
  return cudaGetLastError();
}

}  // namespace hgf
