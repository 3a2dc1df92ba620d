// The lowered form of one network: what the host-side topology compiler
// (pyhgf_b200/topology.py, following pyhgf/utils/get_update_sequence.py:51-351)
// hands to the kernels.  Plain data, passed to the kernel by value as a
// __grid_constant__ parameter (uniform constant-bank reads).
#pragma once
#include <stdint.h>

namespace hgf {

constexpr int MAX_NODES = 32;
constexpr int MAX_EDGES = 96;     // per edge kind
constexpr int MAX_STEPS = 96;     // prediction steps / update steps (summed over the branches of a switching plan)
constexpr int MAX_BRANCHES = 8;   // update-sequence branches of input_custom_sequence
constexpr int MAX_INPUTS = 16;
constexpr int MAX_RECORDS = 640;  // recorded (node, field) leaves
constexpr int MAX_PARAMS = 16;    // batched (per-instance) parameters
constexpr int N_RESP_PARAMS = 4;
constexpr int MAX_PEERS = 8;      // GPUs of one box
constexpr int INTERP_SLOTS = 768;  // attribute-vector capacity of the opcode interpreter

// node_type codes: pyhgf/typing/typing.py:12-52
enum NodeType : int { NODE_CONSTANT = 0, NODE_BINARY = 1, NODE_CONTINUOUS = 2, NODE_VOLATILE = 6 };

// coupling functions: `None` in the reference is LINEAR; the rest is the
// enumerated set a C-ABI can carry (src/math.rs:41-55) plus sin/identity.
enum CouplingFn : int { CF_LINEAR = 0, CF_IDENTITY = 1, CF_RELU = 2, CF_SIGMOID = 3, CF_TANH = 4, CF_LEAKY_RELU = 5, CF_GELU = 6, CF_SIN = 7 };

enum Op : int {
  OP_PRED_BINARY = 1, OP_PRED_CONT = 2, OP_PRED_CONT_MF = 3, OP_PRED_VOLATILE = 4, OP_PRED_VOLATILE_MF = 5,
  OP_PE_BINARY = 10, OP_PE_CONT = 11, OP_PE_VOLATILE = 12,
  OP_POST_CONT = 20, OP_POST_CONT_MF = 21, OP_POST_CONT_EHGF = 22, OP_POST_CONT_EHGF_MF = 23,
  OP_POST_CONT_UNBOUNDED = 24, OP_POST_VOLATILE = 25, OP_POST_VOLATILE_MF = 26,
};

enum VolUpdates : int { VU_UNBOUNDED = 0, VU_EHGF = 1, VU_STANDARD = 2 };

// pyhgf/response.py
enum Response : int {
  RESP_NONE = 0, RESP_FIRST_LEVEL_GAUSSIAN = 1, RESP_TOTAL_GAUSSIAN = 2, RESP_FIRST_LEVEL_BINARY = 3,
  RESP_BINARY_SOFTMAX = 4, RESP_BINARY_SOFTMAX_INVERSE_TEMPERATURE = 5,
};

// Per-node scalar attributes.  slot = node * NF + field; the names are the
// reference's attribute keys (pyhgf/model/add_nodes.py:65-85, 159-184, 330-340)
// with `temp.*` flattened.
enum Field : int {
  F_MEAN = 0, F_EXPECTED_MEAN, F_PRECISION, F_EXPECTED_PRECISION,
  F_TONIC_VOLATILITY, F_TONIC_DRIFT, F_AUTOCONNECTION_STRENGTH, F_OBSERVED,
  F_MEAN_VOL, F_EXPECTED_MEAN_VOL, F_PRECISION_VOL, F_EXPECTED_PRECISION_VOL, F_TONIC_VOLATILITY_VOL,
  F_EFFECTIVE_PRECISION, F_CONDITIONAL_EXPECTED_PRECISION, F_VALUE_PREDICTION_ERROR,
  F_VOLATILITY_PREDICTION_ERROR, F_CURRENT_VARIANCE, F_EFFECTIVE_PRECISION_VOL,
  NF
};

// Slot layout of one instance's attribute vector:
//   [0, n_nodes*NF)                          node fields
//   base_cvp   + off_vp[node]   + k          value_coupling_parents[k]
//   base_cvolp + off_volp[node] + k          volatility_coupling_parents[k]
//   base_cvc   + off_vc[node]   + k          value_coupling_children[k]
//   base_cvolc + off_volc[node] + k          volatility_coupling_children[k]
//   slot_clip                                precision_clipping_value
//   slot_resp + j                            response-function parameter j
struct Program {
  int32_t n_nodes, n_slots;
  int32_t n_pred, n_upd;
  int32_t n_inputs, n_fed;
  int32_t vol_updates, mean_field;
  int32_t base_cvp, base_cvolp, base_cvc, base_cvolc, slot_clip, slot_resp;
  double max_posterior_precision;
  uint8_t node_type[MAX_NODES];
  uint8_t n_vp[MAX_NODES], n_volp[MAX_NODES], n_vc[MAX_NODES], n_volc[MAX_NODES];
  uint8_t off_vp[MAX_NODES], off_volp[MAX_NODES], off_vc[MAX_NODES], off_volc[MAX_NODES];
  uint8_t e_vp[MAX_EDGES], e_vp_fn[MAX_EDGES];  // value parents and the coupling fn the parent applies to this child
  uint8_t e_volp[MAX_EDGES];
  uint8_t e_vc[MAX_EDGES], e_vc_fn[MAX_EDGES];
  uint8_t e_volc[MAX_EDGES];
  uint8_t pred_node[MAX_STEPS], pred_op[MAX_STEPS];
  uint8_t upd_node[MAX_STEPS], upd_op[MAX_STEPS];
  uint8_t input_idx[MAX_INPUTS];
  // input_custom_sequence (network.py:653-767): branch b's steps are pred_*[br_pred[b] .. br_pred[b+1]) and
  // upd_*[br_upd[b] .. br_upd[b+1]); n_branches <= 1 is the ordinary single schedule (n_pred / n_upd)
  int32_t n_branches;
  uint8_t br_pred[MAX_BRANCHES + 1], br_upd[MAX_BRANCHES + 1];
};

// Per-call lists (which leaves are recorded, which slots are per-instance).
// Parameters are kept in launch order: the first n_filter are seeded into the
// attribute vector and carry their tangents through the whole filter; the rest
// are response-function parameters, which cannot influence the beliefs, so their
// tangents exist only inside the response function (Dual<PF + PR> there,
// Dual<PF> in the filter).  param_row[q] is the parameter's row in the caller's
// params / gradient arrays.
struct CallLists {
  int32_t n_rec, n_params, n_filter;
  uint16_t rec_slot[MAX_RECORDS];
  uint16_t param_slot[MAX_PARAMS];
  uint16_t param_row[MAX_PARAMS];
};

// Workspace of the fused cross-GPU reduction, in scalars of width w:
//   result[FUSED_RING][G]   the all-reduced sums of launch (epoch) e live in slot e % FUSED_RING
//   partial[G]              this rank's per-group sums of the running launch
//   ticket (32-bit word, own cache line)   block counter of the running launch
//   done[MAX_PEERS] (32-bit words)         done[r] = number of launches rank r has deposited (written by rank r)
// Protocol (hgf_scan.cuh::fused_deposit): a launch DEPOSITS its sums into every rank's slot and raises its done word
// on every rank, then EXITS -- it never waits for the other ranks' sums.  Whoever consumes the sums waits
// (hgf_fused_reduce_wait: a one-block kernel on the consumer's stream).  Ranks may therefore drift apart by up to
// FUSED_RING / 2 - 1 launches; the only in-kernel wait is that flow control, which a launch passes without spinning
// unless a peer is that far behind.
constexpr int FUSED_RING = 16;
#ifdef __CUDACC__
#define HGF_PROGRAM_HD __host__ __device__
#else
#define HGF_PROGRAM_HD
#endif
HGF_PROGRAM_HD constexpr long long fused_ctl_offset(long long G, long long w) { return ((FUSED_RING + 1) * G * w + 63) / 64 * 64; }
HGF_PROGRAM_HD constexpr long long fused_done_offset(long long G, long long w) { return fused_ctl_offset(G, w) + 64; }
HGF_PROGRAM_HD constexpr long long fused_bytes(long long G, long long w) { return fused_done_offset(G, w) + 64; }

// One launch's arguments (device pointers; see hgf_call_t in include/pyhgf_cuda.h).
struct ScanArgs {
  long long n_instances;
  long long ld;  // leading dimension (row stride, in scalars) of params / gradient / pointwise / trajectories / last_attributes
  int n_steps;
  long long input_group_size, response_group_size;
  const void* input_data;
  const uint8_t* observed;
  const void* time_steps;
  const int32_t* branch_idx;  // [T] branch taken at each step (switching plans only)
  const void* attributes;
  const void* params;
  const void* response_inputs;
  void* trajectories;
  void* last_attributes;
  void* surprise_sum;
  void* surprise_pointwise;
  void* gradient;
  void* reduced;
  const int32_t* group;
  double out_scale;
  int response_fn;
  int tangent_offset;  // this pass differentiates the filter params [tangent_offset, tangent_offset + PF)
  int resp_index;      // ... and the response-function param at this position of the lists (-1: none)
  // register-resident specialisations: which belief leaves (mean, expected_mean, precision,
  // expected_precision per node, + the *_vol twins of volatile nodes, in slot order) are recorded;
  // leaf r is stored in trajectory row popcount(mask below r)
  uint32_t rec_mask[4];
  // Trajectory write-back through TMA (hgf_scan.cuh): when traj_rows > 0 each block stages the step's recorded
  // leaves in shared memory as [traj_rows][SCAN_THREADS] and one thread issues a single 3-D tensor store
  // (box = SCAN_THREADS instances x 1 step x traj_rows leaves) into trajectories[R][T][N]; traj_map is the
  // CUtensorMap of that array (encoded per call by hgf_capi.cu).  traj_rows == 0: plain coalesced stores.
  int traj_rows;
  // generative mode (hgf_call_t::generative): observations are drawn from the input nodes' predictions
  int generative;
  unsigned long long seed;
  // fused cross-GPU reduction (hgf_call_t::peer_workspace, hgf_scan.cuh::fused_all_reduce); world <= 1: off
  int world, rank, n_groups;
  unsigned epoch;
  void* peer_ws[MAX_PEERS];
  void* mc_ws;
  alignas(64) unsigned char traj_map[128];
};

}  // namespace hgf
