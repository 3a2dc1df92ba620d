// Declarations shared by the dispatching translation units (which never see
// kernel bodies) and the instantiating ones, and the pass logic of the gradient.
#pragma once
#include <cuda_runtime.h>

#include "hgf_math.cuh"
#include "hgf_ops.cuh"  // RuntimeTopo, ConstAttributes<>
#include "hgf_program.h"

namespace hgf {

// One kernel launch: the filter runs on T = Dual<B, PF>; PR extra tangent directions live only
// inside the response function.  Defined (and explicitly instantiated) in hgf_launch.cuh users.
template <class T, class Topo, int NSLOTS, int PR>
cudaError_t launch_one(const Program& prog, const CallLists& lists, const ScanArgs& args, cudaStream_t stream);

// Which (PF, PR) kernels of a (topology, base type) were instantiated: bit PF + 5 * PR.
constexpr unsigned variant_bit(int pf, int pr) { return 1u << (pf + 5 * pr); }

// `const_attrs`: no structural attribute (autoconnection strength, tonic drift, couplings, clipping value) is among the
// differentiated parameters of this pass: the register-resident kernels then run their ConstAttributes<> flavour.
template <class B, class Topo, int NSLOTS, unsigned MASK>
cudaError_t launch_variant(int pf, int pr, bool const_attrs, const Program& prog, const CallLists& lists, const ScanArgs& args, cudaStream_t stream) {
#define HGF_TRY(PF_, PR_)                                                                  \
  if constexpr ((MASK & variant_bit(PF_, PR_)) != 0) {                                     \
    if (pf == PF_ && pr == PR_) {                                                          \
      if constexpr (Topo::is_static && PF_ > 0) {                                          \
        if (const_attrs) return launch_one<Dual<B, PF_>, ConstAttributes<Topo>, NSLOTS, PR_>(prog, lists, args, stream); \
      }                                                                                    \
      return launch_one<Dual<B, PF_>, Topo, NSLOTS, PR_>(prog, lists, args, stream);       \
    }                                                                                      \
  }
  HGF_TRY(0, 0) HGF_TRY(1, 0) HGF_TRY(2, 0) HGF_TRY(3, 0) HGF_TRY(4, 0)
  HGF_TRY(0, 1) HGF_TRY(1, 1) HGF_TRY(2, 1) HGF_TRY(3, 1) HGF_TRY(4, 1)
#undef HGF_TRY
  return cudaErrorInvalidValue;
}

// Forward filter: one launch.  Gradient: forward-mode passes of up to four filter-parameter
// tangents (+ one response-parameter tangent) each; the first pass also produces the
// non-gradient outputs.  Returns cudaErrorNotSupported if MASK lacks the kernels needed.
template <class B, class Topo, int NSLOTS, unsigned MASK>
cudaError_t launch_scan(const Program& prog, const CallLists& lists_in, const ScanArgs& args_in, bool want_grad,
                        cudaStream_t stream, int* n_launched) {
  CallLists lists = lists_in;
  ScanArgs a = args_in;
  a.tangent_offset = 0;
  a.resp_index = -1;
  constexpr bool kHasResp = (MASK >> 5) != 0;
  if (!kHasResp || !want_grad) lists.n_filter = lists.n_params;  // response parameters ride along as filter parameters
  if (!want_grad) {
    if ((MASK & 1u) == 0) return cudaErrorNotSupported;
    ++*n_launched;
    return launch_variant<B, Topo, NSLOTS, MASK>(0, 0, false, prog, lists, a, stream);
  }
  if (a.trajectories || a.last_attributes) {
    // trajectories / final attributes of a gradient call: one plain-filter pass (the recording kernels carry no tangents)
    if ((MASK & 1u) == 0) return cudaErrorNotSupported;
    ScanArgs f = a;
    f.surprise_sum = nullptr; f.surprise_pointwise = nullptr; f.gradient = nullptr; f.reduced = nullptr; f.world = 0; f.response_fn = RESP_NONE;
    CallLists fl = lists;
    fl.n_filter = fl.n_params;
    const cudaError_t err = launch_variant<B, Topo, NSLOTS, MASK>(0, 0, false, prog, fl, f, stream);
    ++*n_launched;
    if (err != cudaSuccess) return err;
    a.trajectories = nullptr; a.traj_rows = 0; a.last_attributes = nullptr;
  }
  const int F = lists.n_filter, R = lists.n_params - lists.n_filter;
  // are all differentiated filter parameters beliefs / volatilities / response parameters (hgf_ops.cuh::attr)?
  bool const_attrs = true;
  for (int q = 0; q < F; ++q) {
    const int slot = lists.param_slot[q];
    const bool node_field = slot < prog.n_nodes * NF;
    const int field = slot % NF;
    if (node_field ? (field == F_AUTOCONNECTION_STRENGTH || field == F_TONIC_DRIFT || field == F_OBSERVED) : slot < prog.slot_resp) const_attrs = false;
  }
  int done_f = 0, done_r = 0;
  while (done_f < F || done_r < R) {
    const int remf = F - done_f;
    int pr = done_r < R ? 1 : 0, pf = -1;
    for (int attempt = 0; attempt < 2 && pf < 0; ++attempt) {
      for (int k = 4; k >= 0; --k)  // widest kernel that is not wider than what is left ...
        if ((MASK & variant_bit(k, pr)) && k <= remf && (k > 0 || pr > 0)) { pf = k; break; }
      if (pf < 0)
        for (int k = 1; k <= 4; ++k)  // ... else the narrowest one that covers it
          if ((MASK & variant_bit(k, pr)) && k >= remf) { pf = k; break; }
      if (pf < 0 && pr == 1 && remf > 0) pr = 0; else break;
    }
    if (pf < 0) return cudaErrorNotSupported;
    a.tangent_offset = done_f;
    a.resp_index = pr ? F + done_r : -1;
    const cudaError_t err = launch_variant<B, Topo, NSLOTS, MASK>(pf, pr, const_attrs, prog, lists, a, stream);
    ++*n_launched;
    if (err != cudaSuccess) return err;
    done_f += pf < remf ? pf : remf;
    done_r += pr;
    // later passes only add gradient columns
    a.trajectories = nullptr; a.traj_rows = 0; a.last_attributes = nullptr; a.surprise_sum = nullptr; a.surprise_pointwise = nullptr; a.reduced = nullptr; a.world = 0;
  }
  return cudaSuccess;
}

}  // namespace hgf
