// The consumer's side of the fused cross-GPU reduction (hgf_program.h, hgf_scan.cuh::fused_deposit): the launch that
// produced the per-group sums deposits them on every rank and exits; whoever reads them first enqueues this wait on
// ITS stream.  Declared in include/pyhgf_cuda.h.
#include "../../include/pyhgf_cuda.h"
#include "hgf_scan.cuh"

namespace hgf {
int report_error(const char* msg);
void count_launches(int n);
}  // namespace hgf

extern "C" int hgf_fused_reduce_wait(void* workspace, int32_t n_groups, int32_t dtype, int32_t world_size, uint32_t epoch, void* stream) {
  using namespace hgf;
  if (!workspace || n_groups < 1 || (dtype != 32 && dtype != 64) || world_size < 1 || world_size > MAX_PEERS)
    return report_error("hgf_fused_reduce_wait: invalid argument");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (dtype == 32) fused_wait_kernel<float><<<1, 32, 0, s>>>(workspace, n_groups, world_size, epoch);
  else fused_wait_kernel<double><<<1, 32, 0, s>>>(workspace, n_groups, world_size, epoch);
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return report_error(cudaGetErrorString(err));
  count_launches(1);
  return 0;
}
