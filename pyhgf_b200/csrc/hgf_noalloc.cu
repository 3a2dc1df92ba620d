// hgf_logp_grad_reverse_noalloc (include/pyhgf_cuda.h): the checkpointed reverse sweep of the 3-level binary HGF
// (hgf_reverse.cuh, generated/reverse_binary3.h), float32, for callers that cannot allocate -- code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// hgf_logp_grad's reverse sweep takes its checkpoint array state[T][kCarried][N] from cudaMallocAsync.  Here the array
// lives in a file-scope __device__ array of HGF_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until the
// process exits, so the call makes no cudaMalloc / cudaFree / allocating library call at all.
//   * In HOST code `g_scratch` is only the symbol's host-side stand-in: its device address is asked for once per device
//     with cudaGetSymbolAddress (with that device current) and passed to the kernel as an ordinary pointer.
//   * Every call on a device shares the array.  Nothing in it has to start at zero: a thread reads back only the
//     checkpoints it has itself written earlier in the same launch, so the previous call's contents are never read.
//     Two calls that OVERLAP on the GPU (different streams, or graph branches) race on it: serialise them.
#include <cstdio>
#include <mutex>

#include "../../include/pyhgf_cuda.h"
#include "hgf_reverse.cuh"
#include <generated/reverse_binary3.h>

namespace hgf {
int lower_reverse_call(const hgf_plan_t* plan, const hgf_call_t* c, const char* kernel, const Program** prog, CallLists* lists, ScanArgs* a);
void count_launches(int n);
int report_error(const char* msg);
}  // namespace hgf

using namespace hgf;
using Rev = Rev_binary3<float>;

static_assert(HGF_NOALLOC_SCRATCH_BYTES % 256 == 0, "scratch size");
static_assert(HGF_NOALLOC_MAX_STEP_INSTANCES * Rev::kCarried * sizeof(float) == HGF_NOALLOC_SCRATCH_BYTES,
              "HGF_NOALLOC_MAX_STEP_INSTANCES out of sync with the scratch size");

static __device__ __align__(256) unsigned char g_scratch[HGF_NOALLOC_SCRATCH_BYTES];

// the kernel of hgf_reverse.cuh with the checkpoints in the static scratch (same per-instance code, same launch bound)
template <bool UNIT_DT>
__global__ void __launch_bounds__(SCAN_THREADS, 7)
reverse_binary3_static_scratch_kernel(const __grid_constant__ Program prog, const __grid_constant__ CallLists lists,
                                      const __grid_constant__ ScanArgs args, float* state) {
  long long n = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const bool active = n < args.n_instances;
  if (!active) n = args.n_instances - 1;
  const float total = reverse_instance<float, Rev, UNIT_DT>(prog, lists, args, n, active, state);
  if (args.reduced) {
    const int g = args.group ? __ldg(args.group + n) : 0;
    reduce_into_groups<float>(static_cast<float*>(args.reduced), g, static_cast<float>(args.out_scale) * total, active);
  }
}

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_mutex;
float* g_scratch_address[kMaxDevices] = {};  // device address of g_scratch on each device (nullptr: not asked yet)

int fail(const char* fmt, const char* detail) {
  char buf[512];
  std::snprintf(buf, sizeof buf, fmt, detail);
  return report_error(buf);
}

// Once per device, with that device current: the device address of the scratch array.
int scratch_for_current_device(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail("hgf_logp_grad_reverse_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return fail("hgf_logp_grad_reverse_noalloc: %s", "device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_mutex);
  if (!g_scratch_address[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_scratch);
    if (e != cudaSuccess || !p) return fail("hgf_logp_grad_reverse_noalloc: cudaGetSymbolAddress(scratch): %s", cudaGetErrorString(e));
    g_scratch_address[dev] = static_cast<float*>(p);
  }
  *out = g_scratch_address[dev];
  return 0;
}
}  // namespace

extern "C" int hgf_logp_grad_reverse_noalloc(const hgf_plan_t* plan, const hgf_call_t* call, void* stream_) {
  const Program* prog = nullptr;
  CallLists lists;
  ScanArgs args;
  if (lower_reverse_call(plan, call, "static:binary3", &prog, &lists, &args)) return 1;
  if (call->dtype != 32) return fail("hgf_logp_grad_reverse_noalloc: %s", "float32 only (dtype must be 32)");
  const long long cells = static_cast<long long>(call->n_steps) * call->n_instances;
  if (cells > static_cast<long long>(HGF_NOALLOC_MAX_STEP_INSTANCES)) {
    char buf[256];
    std::snprintf(buf, sizeof buf,
                  "hgf_logp_grad_reverse_noalloc: n_steps x n_instances = %lld exceeds HGF_NOALLOC_MAX_STEP_INSTANCES = %lld "
                  "(the static checkpoint scratch holds %lld bytes); split the batch or call hgf_logp_grad",
                  cells, static_cast<long long>(HGF_NOALLOC_MAX_STEP_INSTANCES), static_cast<long long>(HGF_NOALLOC_SCRATCH_BYTES));
    return report_error(buf);
  }
  if (call->n_instances == 0) return 0;
  float* state = nullptr;
  if (scratch_for_current_device(&state)) return 1;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const unsigned blocks = static_cast<unsigned>((args.n_instances + SCAN_THREADS - 1) / SCAN_THREADS);
  if (args.time_steps) reverse_binary3_static_scratch_kernel<false><<<blocks, SCAN_THREADS, 0, stream>>>(*prog, lists, args, state);
  else reverse_binary3_static_scratch_kernel<true><<<blocks, SCAN_THREADS, 0, stream>>>(*prog, lists, args, state);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail("hgf_logp_grad_reverse_noalloc: kernel launch failed: %s", cudaGetErrorString(e));
  count_launches(1);
  return 0;
}
