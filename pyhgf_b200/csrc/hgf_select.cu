#include <cstdlib>
#include <mutex>
#include <vector>

#include "hgf_dispatch.h"

namespace hgf {

// HGF_FORCE_INTERPRETER=1 (read at plan creation) routes every network through the opcode interpreter:
// used by the tests to run both kernel families on the same inputs.
static std::mutex g_registered_mutex;
static std::vector<const KernelVariant*>& registered() {
  static std::vector<const KernelVariant*> v;
  return v;
}

void register_variant(const KernelVariant* v) {
  std::lock_guard<std::mutex> lock(g_registered_mutex);
  for (const KernelVariant* r : registered())
    if (r == v) return;
  registered().push_back(v);
}

const KernelVariant* select_variant(const Program& prog) {
  const char* force = std::getenv("HGF_FORCE_INTERPRETER");
  if (force && force[0] == '1') return &kInterpreterVariant;
  {
    std::lock_guard<std::mutex> lock(g_registered_mutex);
    for (auto it = registered().rbegin(); it != registered().rend(); ++it)  // newest first
      if ((*it)->matches && (*it)->matches(prog)) return *it;
  }
  for (const KernelVariant* const* v = static_variants(); *v; ++v)
    if ((*v)->matches && (*v)->matches(prog)) return *v;
  return &kInterpreterVariant;
}

// Stream-ordered scratch of the reverse sweep (hgf_reverse.cuh: the checkpoint array).  A pool of the library's own, one
// per device, that KEEPS what it has been given: the device's default pool has a release threshold of 0 and hands the
// array back to the driver at every synchronisation, so that each launch after one would map its 100s of MB again.
cudaError_t scratch_pool(cudaMemPool_t* out) {
  constexpr int kMaxDevices = 64;
  static std::mutex mutex;
  static cudaMemPool_t pools[kMaxDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(mutex);
  if (!pools[dev]) {
    cudaMemPoolProps props = {};
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = dev;
    cudaMemPool_t pool = nullptr;
    e = cudaMemPoolCreate(&pool, &props);
    if (e != cudaSuccess) return e;
    unsigned long long keep = ~0ull;
    e = cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    if (e != cudaSuccess) return e;
    pools[dev] = pool;
  }
  *out = pools[dev];
  return cudaSuccess;
}

}  // namespace hgf
