// Host-side glue of the generated register-resident specialisations:
//   * topo_matches<Topo>  -- does a lowered Program have exactly the topology the
//                            specialisation was generated for?
//   * run_static<...>     -- dispatch to the instantiated kernels of that topology
//                            (falling back to the interpreter for scalar types that
//                            were not instantiated, for record sets outside the
//                            specialisation's belief leaves, or when the final
//                            attribute vector is requested).
#pragma once
#include "hgf_dispatch.h"
#include "hgf_launch_decl.h"
#include "hgf_ops.cuh"

namespace hgf {

template <class Topo>
bool topo_matches(const Program& p) {
  const Topo a(p);
  const RuntimeTopo r(p);
  if (a.n_nodes() != r.n_nodes() || a.n_slots() != r.n_slots() || Topo::kSlots != r.n_slots()) return false;
  if (a.n_pred() != r.n_pred() || a.n_upd() != r.n_upd() || a.n_inputs() != r.n_inputs() || a.n_fed() != r.n_fed()) return false;
  if (a.vol_updates() != r.vol_updates()) return false;
  if (a.slot_clip() != r.slot_clip() || a.slot_resp() != r.slot_resp()) return false;
  for (int i = 0; i < r.n_pred(); ++i)
    if (a.pred_node(i) != r.pred_node(i) || a.pred_op(i) != r.pred_op(i)) return false;
  for (int i = 0; i < r.n_upd(); ++i)
    if (a.upd_node(i) != r.upd_node(i) || a.upd_op(i) != r.upd_op(i)) return false;
  for (int k = 0; k < r.n_inputs(); ++k)
    if (a.input_idx(k) != r.input_idx(k)) return false;
  for (int n = 0; n < r.n_nodes(); ++n) {
    if (a.node_type(n) != r.node_type(n)) return false;
    if (a.n_vp(n) != r.n_vp(n) || a.n_volp(n) != r.n_volp(n) || a.n_vc(n) != r.n_vc(n) || a.n_volc(n) != r.n_volc(n)) return false;
    for (int k = 0; k < r.n_vp(n); ++k)
      if (a.vp(n, k) != r.vp(n, k) || a.vp_fn(n, k) != r.vp_fn(n, k) || a.slot_cvp(n, k) != r.slot_cvp(n, k)) return false;
    for (int k = 0; k < r.n_volp(n); ++k)
      if (a.volp(n, k) != r.volp(n, k) || a.slot_cvolp(n, k) != r.slot_cvolp(n, k)) return false;
    for (int k = 0; k < r.n_vc(n); ++k)
      if (a.vc(n, k) != r.vc(n, k) || a.vc_fn(n, k) != r.vc_fn(n, k) || a.slot_cvc(n, k) != r.slot_cvc(n, k)) return false;
    for (int k = 0; k < r.n_volc(n); ++k)
      if (a.volc(n, k) != r.volc(n, k) || a.slot_cvolc(n, k) != r.slot_cvolc(n, k)) return false;
  }
  return true;
}

// The leaves a specialisation can record: mean, expected_mean, precision,
// expected_precision of every node (+ the *_vol twins of volatile nodes), in
// slot order.  Translates the call's record list into a leaf bit-mask; false if
// a requested slot is not such a leaf or the list is not in ascending order
// (then the row of a leaf would not be its rank in the mask).
template <class Topo>
bool static_record_mask(const Program& p, const CallLists& lists, uint32_t mask[4]) {
  const Topo a(p);
  mask[0] = mask[1] = mask[2] = mask[3] = 0;
  int prev = -1;
  for (int r = 0; r < lists.n_rec; ++r) {
    const int slot = lists.rec_slot[r];
    if (slot <= prev || slot >= a.n_nodes() * NF) return false;
    prev = slot;
    const int node = slot / NF, f = slot % NF;
    int leaf = 0;
    for (int m = 0; m < node; ++m) leaf += (a.node_type(m) == NODE_VOLATILE) ? 8 : 4;
    if (f < 4) leaf += f;
    else if (a.node_type(node) == NODE_VOLATILE && f >= F_MEAN_VOL && f < F_MEAN_VOL + 4) leaf += 4 + (f - F_MEAN_VOL);
    else return false;
    if (leaf >= 128) return false;
    mask[leaf >> 5] |= 1u << (leaf & 31);
  }
  return true;
}

// MASK: the (PF, PR) kernels of this topology that were instantiated (variant_bit in hgf_launch_decl.h).
template <class B, class Topo, unsigned MASK>
cudaError_t run_static(const Program& prog, const CallLists& lists, const ScanArgs& args_in, int dtype, bool want_grad,
                       cudaStream_t stream, int* n_launched) {
  ScanArgs args = args_in;
  const bool rec_ok = !args.last_attributes && (!args.trajectories || static_record_mask<Topo>(prog, lists, args.rec_mask));
  if (rec_ok) {
    const int before = *n_launched;
    const cudaError_t err = launch_scan<B, Topo, Topo::kSlots, MASK>(prog, lists, args, want_grad, stream, n_launched);
    if (err != cudaErrorNotSupported || *n_launched != before) return err;
  }
  return kInterpreterVariant.run(prog, lists, args_in, dtype, want_grad, stream, n_launched);
}

template <class Topo, unsigned MASK32, unsigned MASK64>
cudaError_t run_static_any(const Program& prog, const CallLists& lists, const ScanArgs& args, int dtype, bool want_grad,
                           cudaStream_t stream, int* n_launched) {
  if (dtype == 32) return run_static<float, Topo, MASK32>(prog, lists, args, dtype, want_grad, stream, n_launched);
  return run_static<double, Topo, MASK64>(prog, lists, args, dtype, want_grad, stream, n_launched);
}

}  // namespace hgf
