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
  // Recording kernels (REC) exist for the plain filter only: logp-only and gradient launches run kernels without any
  // recording code, and a gradient call that also records trajectories gets them from an extra plain-filter pass
  // (launch_scan).  That halves the number of gradient kernels to compile.
  constexpr bool kPlain = scalar_traits<T>::n_tangents == 0 && PR == 0;
  const bool rec = args.trajectories != nullptr;
  if constexpr (!kPlain) {
    if (rec) return cudaErrorInvalidValue;
    if constexpr (Topo::is_static) {
      if (!args.time_steps) return go(scan_kernel<T, Topo, NSLOTS, PR, true, false>);
    }
    return go(scan_kernel<T, Topo, NSLOTS, PR, false, false>);
  } else {
    if constexpr (Topo::is_static) {
      if (!args.time_steps)  // dt = 1 folds away (and with it log dt and the per-step load)
        return rec ? go(scan_kernel<T, Topo, NSLOTS, PR, true, true>) : go(scan_kernel<T, Topo, NSLOTS, PR, true, false>);
    }
    return rec ? go(scan_kernel<T, Topo, NSLOTS, PR, false, true>) : go(scan_kernel<T, Topo, NSLOTS, PR, false, false>);
  }
}

}  // namespace hgf
