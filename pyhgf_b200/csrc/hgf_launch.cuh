// Definition of launch_one (instantiated explicitly, one kernel family per
// translation unit so that the build parallelises).
#pragma once
#include "hgf_launch_decl.h"
#include "hgf_scan.cuh"

namespace hgf {

template <class T, class Topo, int NSLOTS, int PR>
cudaError_t launch_one(const Program& prog, const CallLists& lists, const ScanArgs& args, cudaStream_t stream) {
  const long long blocks = (args.n_instances + SCAN_THREADS - 1) / SCAN_THREADS;
  // TMA trajectory write-back: two staging tiles of [traj_rows][SCAN_THREADS] scalars per block
  const size_t smem = args.traj_rows > 0 ? 2 * static_cast<size_t>(args.traj_rows) * SCAN_THREADS * sizeof(base_t<T>) : 0;
  auto go = [&](auto kernel) -> cudaError_t {
    if (smem > 48 * 1024) {
      const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
      if (e != cudaSuccess) return e;
    }
    kernel<<<static_cast<unsigned>(blocks), SCAN_THREADS, smem, stream>>>(prog, lists, args);
    return cudaGetLastError();
  };
  const bool rec = args.trajectories != nullptr;  // logp-only launches run a kernel without any recording code
  if constexpr (Topo::is_static) {
    if (!args.time_steps)  // dt = 1 folds away (and with it log dt and the per-step load)
      return rec ? go(scan_kernel<T, Topo, NSLOTS, PR, true, true>) : go(scan_kernel<T, Topo, NSLOTS, PR, true, false>);
  }
  return rec ? go(scan_kernel<T, Topo, NSLOTS, PR, false, true>) : go(scan_kernel<T, Topo, NSLOTS, PR, false, false>);
}

}  // namespace hgf
