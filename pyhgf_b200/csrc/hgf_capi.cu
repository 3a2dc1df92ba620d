// C ABI of libpyhgf_cuda (include/pyhgf_cuda.h): plan creation (lowering the
// flattened pyhgf network description into the kernels' Program), dispatch to
// the kernel variant, and the host-buffer convenience entry points.
#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include <dlfcn.h>

#include <cuda.h>  // CUtensorMap (types only: the encoder is fetched with cudaGetDriverEntryPoint, no libcuda link)
#include <cstdlib>

#include "../../include/pyhgf_cuda.h"
#include "hgf_dispatch.h"

using namespace hgf;

static_assert(NF == HGF_N_FIELDS, "field count out of sync with the public header");
static_assert(N_RESP_PARAMS == HGF_N_RESPONSE_PARAMETERS, "response parameter count out of sync");
static_assert(FUSED_RING == HGF_FUSED_RING, "fused-reduction ring size out of sync with the public header");

static thread_local std::string g_error;
static std::atomic<long long> g_launches{0};

static int fail(const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_error = buf;
  return 1;
}

namespace hgf {  // shared with the other entry-point translation units (hgf_ef.cu)
void count_launches(int n) { g_launches += n; }
int report_error(const char* msg) { return fail("%s", msg); }
}  // namespace hgf

// gradient_mode auto: the reverse sweep is taken when forward mode would need this many passes (4 filter parameters + 1
// response parameter each).  Measured: the sweep costs 3.6-3.9 plain filters whatever P is, one
// forward-mode pass 1.4-2.4, two passes 3.4-5.4.
constexpr int kReverseFromPasses = 2;

struct hgf_plan {
  Program prog;
  const KernelVariant* variant;
};

extern "C" {

int hgf_abi_version(void) { return HGF_ABI_VERSION; }
const char* hgf_last_error(void) { return g_error.c_str(); }
int64_t hgf_launch_count(void) { return g_launches.load(); }
int64_t hgf_fused_reduce_bytes(int32_t n_groups, int32_t dtype) {
  return (n_groups < 1 || (dtype != 32 && dtype != 64)) ? -1 : fused_bytes(n_groups, dtype / 8);
}
int64_t hgf_fused_reduce_result_offset(int32_t n_groups, int32_t dtype, uint32_t epoch) {
  return (n_groups < 1 || (dtype != 32 && dtype != 64)) ? -1 : static_cast<int64_t>(epoch % FUSED_RING) * n_groups * (dtype / 8);
}

int hgf_plan_create(const hgf_topology_t* t, hgf_plan_t** out) {
  if (!t || !out) return fail("hgf_plan_create: null argument");
  *out = nullptr;
  const int n = t->n_nodes;
  if (n < 1 || n > MAX_NODES) return fail("hgf_plan_create: n_nodes=%d outside [1, %d]", n, MAX_NODES);
  if (t->n_prediction_steps < 0 || t->n_prediction_steps > MAX_STEPS || t->n_update_steps < 0 || t->n_update_steps > MAX_STEPS)
    return fail("hgf_plan_create: schedule longer than %d steps", MAX_STEPS);
  if (t->n_inputs < 0 || t->n_inputs > MAX_INPUTS || t->n_fed < 0 || t->n_fed > t->n_inputs)
    return fail("hgf_plan_create: n_inputs=%d / n_fed=%d invalid (max %d)", t->n_inputs, t->n_fed, MAX_INPUTS);
  if (t->volatility_updates < 0 || t->volatility_updates > 2) return fail("hgf_plan_create: volatility_updates must be 0, 1 or 2");
  hgf_plan* plan = new (std::nothrow) hgf_plan();
  if (!plan) return fail("hgf_plan_create: out of memory");
  Program& p = plan->prog;
  std::memset(&p, 0, sizeof p);
  p.n_nodes = n;
  p.n_pred = t->n_prediction_steps;
  p.n_upd = t->n_update_steps;
  p.n_inputs = t->n_inputs;
  p.n_fed = t->n_fed;
  p.vol_updates = t->volatility_updates;
  p.mean_field = t->mean_field_updates ? 1 : 0;
  p.max_posterior_precision = t->max_posterior_precision;
  p.n_branches = t->n_branches > 1 ? t->n_branches : 0;
  if (p.n_branches > MAX_BRANCHES) { delete plan; return fail("hgf_plan_create: %d branches (max %d)", t->n_branches, MAX_BRANCHES); }
  if (p.n_branches) {
    if (!t->branch_prediction_ptr || !t->branch_update_ptr) { delete plan; return fail("hgf_plan_create: branch pointers are null"); }
    for (int b = 0; b <= p.n_branches; ++b) {
      const int lo = b ? t->branch_prediction_ptr[b - 1] : 0, ulo = b ? t->branch_update_ptr[b - 1] : 0;
      if (t->branch_prediction_ptr[b] < lo || t->branch_prediction_ptr[b] > p.n_pred || t->branch_update_ptr[b] < ulo || t->branch_update_ptr[b] > p.n_upd ||
          (b == 0 && (t->branch_prediction_ptr[0] != 0 || t->branch_update_ptr[0] != 0))) {
        delete plan;
        return fail("hgf_plan_create: branch %d has a bad step range", b);
      }
      p.br_pred[b] = static_cast<uint8_t>(t->branch_prediction_ptr[b]);
      p.br_upd[b] = static_cast<uint8_t>(t->branch_update_ptr[b]);
    }
  }

  auto fill = [&](const int32_t* ptr, const int32_t* idx, uint8_t* count, uint8_t* off, uint8_t* dst, const char* what) -> int {
    if (!ptr) return fail("hgf_plan_create: %s_ptr is null", what);
    if (ptr[n] > MAX_EDGES) return fail("hgf_plan_create: more than %d %s edges", MAX_EDGES, what);
    for (int i = 0; i < n; ++i) {
      const int c = ptr[i + 1] - ptr[i];
      if (c < 0 || c > 255) return fail("hgf_plan_create: bad %s CSR", what);
      count[i] = static_cast<uint8_t>(c);
      off[i] = static_cast<uint8_t>(ptr[i]);
      for (int k = ptr[i]; k < ptr[i + 1]; ++k) {
        if (idx[k] < 0 || idx[k] >= n) return fail("hgf_plan_create: %s index %d out of range", what, idx[k]);
        dst[k] = static_cast<uint8_t>(idx[k]);
      }
    }
    return 0;
  };
  int rc = 0;
  rc = rc || fill(t->value_parents_ptr, t->value_parents_idx, p.n_vp, p.off_vp, p.e_vp, "value_parents");
  rc = rc || fill(t->volatility_parents_ptr, t->volatility_parents_idx, p.n_volp, p.off_volp, p.e_volp, "volatility_parents");
  rc = rc || fill(t->value_children_ptr, t->value_children_idx, p.n_vc, p.off_vc, p.e_vc, "value_children");
  rc = rc || fill(t->volatility_children_ptr, t->volatility_children_idx, p.n_volc, p.off_volc, p.e_volc, "volatility_children");
  if (rc) { delete plan; return 1; }

  for (int i = 0; i < n; ++i) {
    const int ty = t->node_type[i];
    if (ty != NODE_BINARY && ty != NODE_CONTINUOUS && ty != NODE_VOLATILE) {
      delete plan;
      return fail("hgf_plan_create: node %d has type %d; this backend runs binary (1), continuous (2) and volatile (6) state nodes", i, ty);
    }
    p.node_type[i] = static_cast<uint8_t>(ty);
    for (int k = 0; k < p.n_vc[i]; ++k) {
      const int fn = t->value_children_fn ? t->value_children_fn[t->value_children_ptr[i] + k] : 0;
      if (fn < 0 || fn > CF_SIN) { delete plan; return fail("hgf_plan_create: unknown coupling function %d", fn); }
      p.e_vc_fn[p.off_vc[i] + k] = static_cast<uint8_t>(fn);
    }
  }
  // coupling function as seen from the child: the parent's entry for this child
  for (int i = 0; i < n; ++i)
    for (int k = 0; k < p.n_vp[i]; ++k) {
      const int par = p.e_vp[p.off_vp[i] + k];
      int fn = -1;
      for (int j = 0; j < p.n_vc[par]; ++j)
        if (p.e_vc[p.off_vc[par] + j] == i) { fn = p.e_vc_fn[p.off_vc[par] + j]; break; }
      if (fn < 0) { delete plan; return fail("hgf_plan_create: node %d lists value parent %d, which does not list it as a child", i, par); }
      p.e_vp_fn[p.off_vp[i] + k] = static_cast<uint8_t>(fn);
    }
  auto valid_op = [](int op) {
    return (op >= OP_PRED_BINARY && op <= OP_PRED_VOLATILE_MF) || (op >= OP_PE_BINARY && op <= OP_PE_VOLATILE) ||
           (op >= OP_POST_CONT && op <= OP_POST_VOLATILE_MF);
  };
  for (int i = 0; i < p.n_pred; ++i) {
    if (!valid_op(t->prediction_op[i]) || t->prediction_node[i] < 0 || t->prediction_node[i] >= n) { delete plan; return fail("hgf_plan_create: bad prediction step %d", i); }
    p.pred_node[i] = static_cast<uint8_t>(t->prediction_node[i]);
    p.pred_op[i] = static_cast<uint8_t>(t->prediction_op[i]);
  }
  for (int i = 0; i < p.n_upd; ++i) {
    if (!valid_op(t->update_op[i]) || t->update_node[i] < 0 || t->update_node[i] >= n) { delete plan; return fail("hgf_plan_create: bad update step %d", i); }
    if (t->update_op[i] == OP_POST_CONT_UNBOUNDED && p.n_volc[t->update_node[i]] == 0) { delete plan; return fail("hgf_plan_create: unbounded update on node %d without volatility children", t->update_node[i]); }
    p.upd_node[i] = static_cast<uint8_t>(t->update_node[i]);
    p.upd_op[i] = static_cast<uint8_t>(t->update_op[i]);
  }
  for (int k = 0; k < p.n_inputs; ++k) {
    if (t->input_idxs[k] < 0 || t->input_idxs[k] >= n) { delete plan; return fail("hgf_plan_create: input index out of range"); }
    p.input_idx[k] = static_cast<uint8_t>(t->input_idxs[k]);
  }
  int off = n * NF;
  p.base_cvp = off; off += t->value_parents_ptr[n];
  p.base_cvolp = off; off += t->volatility_parents_ptr[n];
  p.base_cvc = off; off += t->value_children_ptr[n];
  p.base_cvolc = off; off += t->volatility_children_ptr[n];
  p.slot_clip = off++;
  p.slot_resp = off; off += N_RESP_PARAMS;
  p.n_slots = off;
  plan->variant = p.n_branches ? &kInterpreterVariant : select_variant(p);  // switching needs the rolled schedule
  if (!plan->variant->matches && p.n_slots > INTERP_SLOTS) {
    delete plan;
    return fail("hgf_plan_create: network needs %d attribute slots; the interpreter holds %d", off, INTERP_SLOTS);
  }
  *out = plan;
  return 0;
}

void hgf_plan_destroy(hgf_plan_t* plan) { delete plan; }

int hgf_register_specialisation(const char* path) {
  if (!path) return fail("hgf_register_specialisation: null path");
  void* so = dlopen(path, RTLD_NOW | RTLD_LOCAL);
  if (!so) return fail("hgf_register_specialisation: %s", dlerror());
  using Getter = const void* (*)();
  Getter get = reinterpret_cast<Getter>(dlsym(so, "hgf_specialisation_variant"));
  auto abi = reinterpret_cast<int (*)()>(dlsym(so, "hgf_specialisation_abi"));
  if (!get || !abi) return fail("hgf_register_specialisation: %s is not a pyhgf_b200 specialisation", path);
  if (abi() != HGF_ABI_VERSION) return fail("hgf_register_specialisation: %s was built for ABI %d, this library is %d", path, abi(), HGF_ABI_VERSION);
  register_variant(static_cast<const KernelVariant*>(get()));
  return 0;
}

int32_t hgf_plan_n_slots(const hgf_plan_t* plan) { return plan ? plan->prog.n_slots : -1; }
int32_t hgf_plan_slot_coupling(const hgf_plan_t* plan, int32_t kind, int32_t node, int32_t pos) {
  if (!plan || node < 0 || node >= plan->prog.n_nodes || pos < 0) return -1;
  const Program& p = plan->prog;
  switch (kind) {
    case 0: return pos < p.n_vp[node] ? p.base_cvp + p.off_vp[node] + pos : -1;
    case 1: return pos < p.n_volp[node] ? p.base_cvolp + p.off_volp[node] + pos : -1;
    case 2: return pos < p.n_vc[node] ? p.base_cvc + p.off_vc[node] + pos : -1;
    case 3: return pos < p.n_volc[node] ? p.base_cvolc + p.off_volc[node] + pos : -1;
  }
  return -1;
}
int32_t hgf_plan_slot_clipping(const hgf_plan_t* plan) { return plan ? plan->prog.slot_clip : -1; }
int32_t hgf_plan_slot_response(const hgf_plan_t* plan) { return plan ? plan->prog.slot_resp : -1; }
const char* hgf_plan_kernel_name(const hgf_plan_t* plan) { return plan ? plan->variant->name : ""; }

}  // extern "C"

// ---------------------------------------------------------------------------
// launch
// ---------------------------------------------------------------------------
static int check_call(const hgf_plan_t* plan, const hgf_call_t* c, bool want_grad, CallLists& lists) {
  if (!plan || !c) return fail("null plan or call");
  const Program& p = plan->prog;
  if (c->dtype != 32 && c->dtype != 64) return fail("dtype must be 32 or 64, got %d", c->dtype);
  if (c->n_instances < 0 || c->n_steps < 0) return fail("negative n_instances / n_steps");
  if (!c->attributes) return fail("attributes is null");
  if (p.n_fed > 0 && c->n_steps > 0 && !c->input_data && !c->generative) return fail("input_data is null");
  if (c->generative && want_grad) return fail("generative mode runs through hgf_filter only");
  if (c->generative && !(c->trajectories && c->n_records > 0)) return fail("generative mode needs recorded trajectories (the rollouts are the output)");
  if (c->n_params < 0 || c->n_params > MAX_PARAMS) return fail("n_params=%d outside [0, %d]", c->n_params, MAX_PARAMS);
  if (c->n_params > 0 && (!c->param_slots || (!c->params && c->n_instances > 0))) return fail("params / param_slots is null");
  if (c->n_records < 0 || c->n_records > MAX_RECORDS) return fail("n_records=%d outside [0, %d]", c->n_records, MAX_RECORDS);
  if (c->trajectories && c->n_records > 0 && !c->record_slots) return fail("record_slots is null");
  if (c->response_fn < RESP_NONE || c->response_fn > RESP_BINARY_SOFTMAX_INVERSE_TEMPERATURE) return fail("unknown response_fn %d", c->response_fn);
  if ((c->response_fn == RESP_BINARY_SOFTMAX || c->response_fn == RESP_BINARY_SOFTMAX_INVERSE_TEMPERATURE) && !c->response_inputs)
    return fail("response function %d needs response_inputs", c->response_fn);
  if ((c->response_fn == RESP_FIRST_LEVEL_GAUSSIAN) && p.n_nodes < 2) return fail("first_level_gaussian_surprise needs a node 1");
  if (c->input_group_size < 0 || c->response_group_size < 0) return fail("negative group size");
  if (p.n_branches && c->n_steps > 0 && !c->branch_idx) return fail("a switching plan needs branch_idx");
  if (!p.n_branches && c->branch_idx) return fail("branch_idx given, but the plan has a single update sequence");
  if (p.n_branches && c->generative) return fail("generative mode does not take a switching plan");
  if (c->world_size > 1) {
    if (c->world_size > MAX_PEERS) return fail("world_size=%d; the fused reduction spans at most %d GPUs", c->world_size, MAX_PEERS);
    if (c->rank < 0 || c->rank >= c->world_size) return fail("rank %d outside [0, %d)", c->rank, c->world_size);
    if (!c->peer_workspace || c->n_groups < 1) return fail("fused reduction needs peer_workspace and n_groups >= 1");
    for (int p = 0; p < c->world_size; ++p)
      if (!c->peer_workspace[p]) return fail("peer_workspace[%d] is null", p);
    if (c->reduced) return fail("fused reduction writes into the workspace: reduced must be null");
  }
  if (want_grad && !c->gradient) return fail("hgf_logp_grad: gradient is null");
  if (want_grad && c->n_params < 1) return fail("hgf_logp_grad: no per-instance parameter to differentiate");
  if (!want_grad && c->gradient) return fail("hgf_filter: gradient must be null (use hgf_logp_grad)");
  std::memset(&lists, 0, sizeof lists);
  lists.n_params = c->n_params;
  // launch order: parameters of the filter first, response-function parameters last (their tangents
  // never enter the filter: hgf_program.h); param_row keeps each one's row in the caller's arrays
  int k = 0;
  for (int pass = 0; pass < 2; ++pass)
    for (int i = 0; i < c->n_params; ++i) {
      if (c->param_slots[i] < 0 || c->param_slots[i] >= p.n_slots) return fail("param slot %d out of range", c->param_slots[i]);
      const bool is_response = c->param_slots[i] >= p.slot_resp;
      if (is_response != (pass == 1)) continue;
      lists.param_slot[k] = static_cast<uint16_t>(c->param_slots[i]);
      lists.param_row[k] = static_cast<uint16_t>(i);
      ++k;
      if (pass == 0) lists.n_filter = k;
    }
  lists.n_rec = c->trajectories ? c->n_records : 0;
  for (int i = 0; i < lists.n_rec; ++i) {
    if (c->record_slots[i] < 0 || c->record_slots[i] >= p.n_slots) return fail("record slot %d out of range", c->record_slots[i]);
    lists.rec_slot[i] = static_cast<uint16_t>(c->record_slots[i]);
  }
  return 0;
}

// Tensor map of trajectories[R][T][N] for the kernels' TMA write-back (ScanArgs::traj_map): dims (N, T, R),
// box (SCAN_THREADS instances, 1 step, R leaves).  Returns false when the array does not meet TMA's
// alignment rules or the staging tile would not fit -- the kernel then uses plain coalesced stores.
static bool encode_trajectory_map(ScanArgs& a, int n_rec, int width) {
  const char* env = std::getenv("HGF_NO_TMA");  // read per call: the tests run both write-back paths in one process
  const bool disabled = env && env[0] == '1';
  constexpr int kThreads = 64;  // SCAN_THREADS (hgf_scan.cuh)
  if (disabled || n_rec < 1 || n_rec > 256) return false;
  if (2ull * n_rec * kThreads * width > 96 * 1024) return false;
  if ((a.n_instances * width) % 16 != 0 || reinterpret_cast<uintptr_t>(a.trajectories) % 16 != 0) return false;
  if (a.n_instances > 0x7fffffffLL || a.n_steps < 1) return false;
  using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn encode = [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
      fn = nullptr;
    return reinterpret_cast<EncodeFn>(fn);
  }();
  if (!encode) return false;
  static_assert(sizeof(CUtensorMap) == sizeof(a.traj_map), "CUtensorMap size");
  const cuuint64_t dims[3] = {static_cast<cuuint64_t>(a.n_instances), static_cast<cuuint64_t>(a.n_steps), static_cast<cuuint64_t>(n_rec)};
  const cuuint64_t strides[2] = {static_cast<cuuint64_t>(a.n_instances) * width,
                                 static_cast<cuuint64_t>(a.n_instances) * a.n_steps * width};
  const cuuint32_t box[3] = {kThreads, 1, static_cast<cuuint32_t>(n_rec)};
  const cuuint32_t elem[3] = {1, 1, 1};
  const CUresult r = encode(reinterpret_cast<CUtensorMap*>(a.traj_map), width == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT64,
                            3, a.trajectories, dims, strides, box, elem, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                            CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return false;
  a.traj_rows = n_rec;
  return true;
}

static int launch(const hgf_plan_t* plan, const hgf_call_t* c, cudaStream_t stream, bool want_grad, long long ld = 0) {
  CallLists lists;
  if (check_call(plan, c, want_grad, lists)) return 1;
  if (c->n_instances == 0) return 0;
  ScanArgs a;
  std::memset(&a, 0, sizeof a);
  a.n_instances = c->n_instances;
  a.ld = ld > 0 ? ld : c->n_instances;
  a.n_steps = c->n_steps;
  a.input_group_size = c->input_group_size;
  a.response_group_size = c->response_group_size;
  a.input_data = c->input_data;
  a.observed = c->observed;
  a.time_steps = c->time_steps;
  a.branch_idx = plan->prog.n_branches ? c->branch_idx : nullptr;
  a.attributes = c->attributes;
  a.params = c->params;
  a.response_inputs = c->response_inputs;
  a.trajectories = c->trajectories;
  a.last_attributes = c->last_attributes;
  a.surprise_sum = c->surprise_sum;
  a.surprise_pointwise = c->surprise_pointwise;
  a.gradient = want_grad ? c->gradient : nullptr;
  a.reduced = c->reduced;
  a.group = c->group;
  a.out_scale = c->out_scale;
  a.response_fn = c->response_fn;
  a.generative = c->generative ? 1 : 0;
  a.seed = c->seed;
  if (c->world_size > 1) {
    a.world = c->world_size;
    a.rank = c->rank;
    a.n_groups = c->n_groups;
    a.epoch = c->epoch;
    for (int p = 0; p < c->world_size; ++p) a.peer_ws[p] = c->peer_workspace[p];
    a.mc_ws = c->multicast_workspace;
  }
  if (a.trajectories && lists.n_rec > 0) encode_trajectory_map(a, lists.n_rec, c->dtype == 32 ? 4 : 8);
  int n_launched = 0;
  // the reverse sweep (hgf_call_t::gradient_mode): a plain logp + gradient call on a topology that has one
  if (want_grad && c->gradient_mode != 1) {
    const bool plain = !c->trajectories && !c->surprise_pointwise && !c->last_attributes && c->world_size <= 1 && !c->generative &&
                       !plan->prog.n_branches && c->response_fn != RESP_NONE && plan->prog.n_fed == plan->prog.n_inputs;
    const bool available = plan->variant->reverse != nullptr && plain;
    if (c->gradient_mode == 2 && !available)
      return fail("gradient_mode = reverse: %s", plan->variant->reverse ? "the call asks for outputs only the forward-mode kernels produce"
                                                                          : "this plan's topology has no generated reverse sweep");
    const int n_resp = lists.n_params - lists.n_filter;
    const int forward_passes = std::max((lists.n_filter + 3) / 4, n_resp);
    // no register-resident gradient kernel in this variant (e.g. the 16-node network): forward mode would be the interpreter's
    const unsigned mask = c->dtype == 32 ? plan->variant->mask32 : plan->variant->mask64;
    const bool native_forward = (mask & ~1u) != 0;
    if (available && (c->gradient_mode == 2 || forward_passes >= kReverseFromPasses || !native_forward)) {
      lists.n_filter = lists.n_params;
      const cudaError_t rerr = plan->variant->reverse(plan->prog, lists, a, c->dtype, stream, &n_launched);
      g_launches += n_launched;
      if (rerr != cudaSuccess) return fail("reverse sweep launch failed: %s", cudaGetErrorString(rerr));
      return 0;
    }
  }
  cudaError_t err = run_variant(*plan->variant, plan->prog, lists, a, c->dtype, want_grad, stream, &n_launched);
  g_launches += n_launched;
  if (err != cudaSuccess) return fail("kernel launch failed: %s", cudaGetErrorString(err));
  return 0;
}

namespace hgf {
// For hgf_noalloc.cu (hgf_logp_grad_reverse_noalloc): validate a plain logp + gradient call on the plan of the topology
// `kernel` ("static:binary3") and lower it to the reverse sweep's arguments.  Launches and allocates nothing.
int lower_reverse_call(const hgf_plan_t* plan, const hgf_call_t* c, const char* kernel, const Program** prog, CallLists* lists, ScanArgs* a) {
  if (check_call(plan, c, true, *lists)) return 1;
  if (std::strcmp(plan->variant->name, kernel) != 0)
    return fail("this entry point runs the '%s' kernel only; the plan dispatches to '%s'", kernel, plan->variant->name);
  if (c->trajectories || c->surprise_pointwise || c->last_attributes || c->world_size > 1 || c->generative || plan->prog.n_branches ||
      c->response_fn == RESP_NONE || c->response_fn == RESP_TOTAL_GAUSSIAN)
    return fail("the reverse sweep returns the summed surprise and its gradient only (no trajectories / pointwise output / fused "
                "cross-GPU reduction) and needs a per-trial response function");
  lists->n_filter = lists->n_params;
  std::memset(a, 0, sizeof *a);
  a->n_instances = c->n_instances;
  a->ld = c->n_instances;
  a->n_steps = c->n_steps;
  a->input_group_size = c->input_group_size;
  a->response_group_size = c->response_group_size;
  a->input_data = c->input_data;
  a->observed = c->observed;
  a->time_steps = c->time_steps;
  a->attributes = c->attributes;
  a->params = c->params;
  a->response_inputs = c->response_inputs;
  a->surprise_sum = c->surprise_sum;
  a->gradient = c->gradient;
  a->reduced = c->reduced;
  a->group = c->group;
  a->out_scale = c->out_scale;
  a->response_fn = c->response_fn;
  *prog = &plan->prog;
  return 0;
}
}  // namespace hgf

extern "C" int hgf_filter(const hgf_plan_t* plan, const hgf_call_t* call, void* stream) {
  return launch(plan, call, static_cast<cudaStream_t>(stream), false);
}
extern "C" int hgf_logp_grad(const hgf_plan_t* plan, const hgf_call_t* call, void* stream) {
  return launch(plan, call, static_cast<cudaStream_t>(stream), true);
}

// ---------------------------------------------------------------------------
// host-buffer entry points: stage in, run, stage out
// ---------------------------------------------------------------------------
namespace {
// Per-thread, per-device staging context of the *_host entry points: one non-blocking stream and one device
// arena that only grows, so that a sampler calling hgf_logp_grad_host thousands of times pays for stream
// creation and device allocation once (they cost more than the copies themselves for a 65 536-instance batch).
constexpr int kHostChunks = 8;         // most slices of the chunk pipeline (streams are created once)
constexpr int kDefaultHostChunks = 4;  // HGF_HOST_CHUNKS=n overrides (0 / 1: one-shot path)
struct HostContext {
  int device = -1;
  cudaStream_t stream = nullptr;
  cudaStream_t side[kHostChunks - 1] = {};  // chunk pipeline of the *_host entry points
  cudaEvent_t shared_ready = nullptr, chunk_done[kHostChunks - 1] = {};
  char* arena = nullptr;
  size_t capacity = 0;
  ~HostContext() { release(); }
  void release() {
    if (arena) cudaFree(arena);
    if (stream) cudaStreamDestroy(stream);
    for (auto& s : side) { if (s) cudaStreamDestroy(s); s = nullptr; }
    for (auto& e : chunk_done) { if (e) cudaEventDestroy(e); e = nullptr; }
    if (shared_ready) cudaEventDestroy(shared_ready);
    shared_ready = nullptr;
    arena = nullptr; stream = nullptr; capacity = 0; device = -1;
  }
  cudaError_t prepare(int dev, size_t bytes) {
    cudaError_t e = cudaSetDevice(dev);
    if (e != cudaSuccess) return e;
    if (dev != device) { release(); device = dev; }
    if (!stream) {
      e = cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking);
      if (e != cudaSuccess) return e;
      for (auto& s : side)
        if ((e = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking)) != cudaSuccess) return e;
      for (auto& ev : chunk_done)
        if ((e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)) != cudaSuccess) return e;
      if ((e = cudaEventCreateWithFlags(&shared_ready, cudaEventDisableTiming)) != cudaSuccess) return e;
    }
    if (bytes > capacity) {
      if (arena) { cudaStreamSynchronize(stream); cudaFree(arena); arena = nullptr; capacity = 0; }
      const size_t want = bytes + bytes / 4;
      e = cudaMalloc(reinterpret_cast<void**>(&arena), want);
      if (e != cudaSuccess) return e;
      capacity = want;
    }
    return cudaSuccess;
  }
};
thread_local HostContext g_host;

constexpr size_t kAlign = 256;
inline size_t aligned(size_t b) { return (b + kAlign - 1) / kAlign * kAlign; }

struct Staging {
  HostContext* ctx;
  size_t offset = 0;
  struct Out { void* dev; void* host; size_t bytes; };
  std::vector<Out> outs;
  cudaError_t err = cudaSuccess;
  void* take(size_t bytes) {
    void* d = ctx->arena + offset;
    offset += aligned(bytes);
    return d;
  }
  const void* in(const void* host, size_t bytes) {
    if (!host || bytes == 0 || err != cudaSuccess) return nullptr;
    void* d = take(bytes);
    err = cudaMemcpyAsync(d, host, bytes, cudaMemcpyHostToDevice, ctx->stream);
    return d;
  }
  void* out(void* host, size_t bytes, bool init_from_host = false) {
    if (!host || bytes == 0 || err != cudaSuccess) return nullptr;
    void* d = take(bytes);
    if (init_from_host) err = cudaMemcpyAsync(d, host, bytes, cudaMemcpyHostToDevice, ctx->stream);
    outs.push_back({d, host, bytes});
    return d;
  }
  void finish() {
    if (err == cudaSuccess)
      for (auto& o : outs) {
        err = cudaMemcpyAsync(o.host, o.dev, o.bytes, cudaMemcpyDeviceToHost, ctx->stream);
        if (err != cudaSuccess) break;
      }
    cudaError_t e2 = cudaStreamSynchronize(ctx->stream);
    if (err == cudaSuccess) err = e2;
  }
};
}  // namespace

// Large logp(+grad) batches: the instance range is cut into kHostChunks slices (whole series groups, multiples of
// the block size), each with its own stream: slice k+1's host->device copies and slice k-1's device->host copies run
// under slice k's kernel.  The device arrays keep the full [.][N] layout (row stride = N: ScanArgs::ld), so a slice is
// a pointer offset.  Returns -1 when the call is not eligible (then the one-shot path below runs), else the result.
static int run_host_chunked(const hgf_plan_t* plan, const hgf_call_t* c, bool want_grad, int n_groups) {
  const char* env = std::getenv("HGF_HOST_CHUNKS");
  int n_chunks = kDefaultHostChunks;
  if (env && env[0] >= '0' && env[0] <= '9') n_chunks = env[0] - '0';
  if (n_chunks < 2) return -1;
  if (n_chunks > kHostChunks) n_chunks = kHostChunks;
  const Program& p = plan->prog;
  const long long N = c->n_instances;
  if (N < 32768 || c->trajectories || c->surprise_pointwise || c->last_attributes || c->world_size > 1 || c->generative || c->branch_idx) return -1;
  if (!c->surprise_sum && !c->gradient) return -1;
  auto lcm = [](long long a, long long b) { long long x = a, y = b; while (y) { long long t = x % y; x = y; y = t; } return a / x * b; };
  long long unit = 64;
  if (c->input_group_size > 0) unit = lcm(unit, c->input_group_size);
  if (c->response_group_size > 0) unit = lcm(unit, c->response_group_size);
  const long long per = ((N + n_chunks - 1) / n_chunks + unit - 1) / unit * unit;
  if (per <= 0 || per >= N) return -1;
  const size_t w = c->dtype == 32 ? 4 : 8, T = static_cast<size_t>(c->n_steps), F = static_cast<size_t>(p.n_fed), P = static_cast<size_t>(c->n_params);
  const size_t in_series = c->input_group_size > 0 ? (N + c->input_group_size - 1) / c->input_group_size : 1;
  const size_t resp_series = c->response_group_size > 0 ? (N + c->response_group_size - 1) / c->response_group_size : 1;
  HostContext& h = g_host;
  size_t off = 0;
  auto take = [&](const void* host, size_t bytes) -> char* {
    if (!host || !bytes) return nullptr;
    char* d = h.arena + off;
    off += aligned(bytes);
    return d;
  };
  char* d_in = take(c->input_data, in_series * T * F * w);
  char* d_obs = take(c->observed, in_series * T * F);
  char* d_ts = take(c->time_steps, T * w);
  char* d_attr = take(c->attributes, static_cast<size_t>(p.n_slots) * w);
  char* d_par = take(c->params, P * N * w);
  char* d_resp = take(c->response_inputs, resp_series * T * w);
  char* d_group = take(c->group, N * sizeof(int32_t));
  char* d_sum = take(c->surprise_sum, N * w);
  char* d_grad = want_grad ? take(c->gradient, P * N * w) : nullptr;
  char* d_red = take(c->reduced, static_cast<size_t>(n_groups) * w);
  if (off > h.capacity) return -1;  // (prepare() sized the arena with the same terms; defensive)
  cudaError_t err = cudaSuccess;
  auto ok = [&](cudaError_t e) { if (err == cudaSuccess) err = e; };
  auto h2d = [&](char* dst, const void* src, size_t bytes, cudaStream_t s) { if (dst && bytes) ok(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, s)); };
  h2d(d_ts, c->time_steps, T * w, h.stream);
  h2d(d_attr, c->attributes, static_cast<size_t>(p.n_slots) * w, h.stream);
  h2d(d_red, c->reduced, static_cast<size_t>(n_groups) * w, h.stream);
  if (in_series == 1) { h2d(d_in, c->input_data, T * F * w, h.stream); h2d(d_obs, c->observed, T * F, h.stream); }
  if (resp_series == 1) h2d(d_resp, c->response_inputs, T * w, h.stream);
  ok(cudaEventRecord(h.shared_ready, h.stream));
  int rc = 0, k = 0;
  for (long long n0 = 0; n0 < N && rc == 0 && err == cudaSuccess; n0 += per, ++k) {
    const long long nc = (n0 + per < N) ? per : N - n0;
    cudaStream_t s = k == 0 ? h.stream : h.side[k - 1];
    if (k > 0) ok(cudaStreamWaitEvent(s, h.shared_ready, 0));
    hgf_call_t d = *c;
    d.n_instances = nc;
    if (in_series > 1) {
      const size_t g0 = n0 / c->input_group_size, g1 = (n0 + nc + c->input_group_size - 1) / c->input_group_size;
      h2d(d_in + g0 * T * F * w, static_cast<const char*>(c->input_data) + g0 * T * F * w, (g1 - g0) * T * F * w, s);
      if (d_obs) h2d(d_obs + g0 * T * F, c->observed + g0 * T * F, (g1 - g0) * T * F, s);
      d.input_data = d_in + g0 * T * F * w;
      d.observed = d_obs ? reinterpret_cast<const uint8_t*>(d_obs + g0 * T * F) : nullptr;
    } else {
      d.input_data = d_in;
      d.observed = reinterpret_cast<const uint8_t*>(d_obs);
    }
    if (resp_series > 1) {
      const size_t g0 = n0 / c->response_group_size, g1 = (n0 + nc + c->response_group_size - 1) / c->response_group_size;
      h2d(d_resp + g0 * T * w, static_cast<const char*>(c->response_inputs) + g0 * T * w, (g1 - g0) * T * w, s);
      d.response_inputs = d_resp + g0 * T * w;
    } else {
      d.response_inputs = d_resp;
    }
    if (d_par) ok(cudaMemcpy2DAsync(d_par + n0 * w, N * w, static_cast<const char*>(c->params) + n0 * w, N * w, nc * w, P, cudaMemcpyHostToDevice, s));
    if (d_group) h2d(d_group + n0 * sizeof(int32_t), c->group + n0, nc * sizeof(int32_t), s);
    d.time_steps = d_ts;
    d.attributes = d_attr;
    d.params = d_par ? d_par + n0 * w : nullptr;
    d.group = d_group ? reinterpret_cast<const int32_t*>(d_group + n0 * sizeof(int32_t)) : nullptr;
    d.surprise_sum = d_sum ? d_sum + n0 * w : nullptr;
    d.gradient = d_grad ? d_grad + n0 * w : nullptr;
    d.reduced = d_red;
    if (err == cudaSuccess) rc = launch(plan, &d, s, want_grad, N);
    if (rc == 0) {
      if (d_sum) ok(cudaMemcpyAsync(static_cast<char*>(c->surprise_sum) + n0 * w, d_sum + n0 * w, nc * w, cudaMemcpyDeviceToHost, s));
      if (d_grad) ok(cudaMemcpy2DAsync(static_cast<char*>(c->gradient) + n0 * w, N * w, d_grad + n0 * w, N * w, nc * w, P, cudaMemcpyDeviceToHost, s));
    }
    if (k > 0) { ok(cudaEventRecord(h.chunk_done[k - 1], s)); ok(cudaStreamWaitEvent(h.stream, h.chunk_done[k - 1], 0)); }
  }
  if (rc == 0 && err == cudaSuccess && d_red) ok(cudaMemcpyAsync(c->reduced, d_red, static_cast<size_t>(n_groups) * w, cudaMemcpyDeviceToHost, h.stream));
  const cudaError_t e2 = cudaStreamSynchronize(h.stream);
  for (auto& s : h.side) cudaStreamSynchronize(s);
  if (rc) return rc;
  ok(e2);
  if (err != cudaSuccess) return fail("host staging (chunked) failed: %s", cudaGetErrorString(err));
  return 0;
}

static int run_host(const hgf_plan_t* plan, const hgf_call_t* c_in, int device, bool want_grad) {
  CallLists lists;
  // With a fused cross-GPU reduction the sums live in the workspace; the host entry points accept a HOST `reduced`
  // all the same and deliver the ALL-REDUCED sums there: launch, wait for every rank's deposit, copy the slot out.
  hgf_call_t c_local = *c_in;
  void* host_all_reduced = nullptr;
  if (c_in->world_size > 1 && c_in->reduced) { host_all_reduced = c_in->reduced; c_local.reduced = nullptr; }
  const hgf_call_t* c = &c_local;
  if (check_call(plan, c, want_grad, lists)) return 1;
  const Program& p = plan->prog;
  const size_t w = c->dtype == 32 ? 4 : 8;
  const size_t N = static_cast<size_t>(c->n_instances), T = static_cast<size_t>(c->n_steps);
  const size_t n_in_series = c->input_group_size > 0 ? (N + c->input_group_size - 1) / c->input_group_size : 1;
  const size_t n_resp_series = c->response_group_size > 0 ? (N + c->response_group_size - 1) / c->response_group_size : 1;
  int n_groups = 1;
  if (c->reduced && c->group) {
    if (c->n_groups > 0) n_groups = c->n_groups;  // the caller's count saves a pass over group[]
    else
      for (size_t i = 0; i < N; ++i) n_groups = c->group[i] + 1 > n_groups ? c->group[i] + 1 : n_groups;
  }
  // (pointer, bytes) of every staged array, in the order they are taken below
  const size_t sz_in = n_in_series * T * p.n_fed * w, sz_obs = n_in_series * T * p.n_fed, sz_ts = T * w;
  const size_t sz_attr = static_cast<size_t>(p.n_slots) * w, sz_par = static_cast<size_t>(c->n_params) * N * w;
  const size_t sz_resp = n_resp_series * T * w, sz_group = N * sizeof(int32_t);
  const size_t sz_traj = static_cast<size_t>(lists.n_rec) * T * N * w, sz_last = static_cast<size_t>(p.n_nodes) * NF * N * w;
  const size_t sz_sum = N * w, sz_pw = T * N * w, sz_grad = sz_par, sz_red = static_cast<size_t>(n_groups) * w;
  auto need = [](const void* ptr, size_t b) { return ptr && b ? aligned(b) : size_t(0); };
  const size_t sz_br = T * sizeof(int32_t);
  const size_t total = need(c->branch_idx, sz_br) + need(c->input_data, sz_in) + need(c->observed, sz_obs) + need(c->time_steps, sz_ts) + need(c->attributes, sz_attr) +
                       need(c->params, sz_par) + need(c->response_inputs, sz_resp) + need(c->group, sz_group) + need(c->trajectories, sz_traj) +
                       need(c->last_attributes, sz_last) + need(c->surprise_sum, sz_sum) + need(c->surprise_pointwise, sz_pw) +
                       (want_grad ? need(c->gradient, sz_grad) : 0) + need(c->reduced, sz_red);
  cudaError_t e = g_host.prepare(device, total);
  if (e != cudaSuccess) return fail("host staging on device %d (%zu bytes): %s", device, total, cudaGetErrorString(e));
  if (int rc = run_host_chunked(plan, c, want_grad, n_groups); rc >= 0) return rc;
  Staging st;
  st.ctx = &g_host;
  hgf_call_t d = *c;
  d.input_data = st.in(c->input_data, sz_in);
  d.observed = static_cast<const uint8_t*>(st.in(c->observed, sz_obs));
  d.time_steps = st.in(c->time_steps, sz_ts);
  d.branch_idx = static_cast<const int32_t*>(st.in(c->branch_idx, sz_br));
  d.attributes = st.in(c->attributes, sz_attr);
  d.params = st.in(c->params, sz_par);
  d.response_inputs = st.in(c->response_inputs, sz_resp);
  d.group = static_cast<const int32_t*>(st.in(c->group, sz_group));
  d.trajectories = st.out(c->trajectories, sz_traj);
  d.last_attributes = st.out(c->last_attributes, sz_last);
  d.surprise_sum = st.out(c->surprise_sum, sz_sum);
  d.surprise_pointwise = st.out(c->surprise_pointwise, sz_pw);
  d.gradient = want_grad ? st.out(c->gradient, sz_grad) : nullptr;
  d.reduced = st.out(c->reduced, sz_red, true);
  int rc = 0;
  if (st.err == cudaSuccess) rc = launch(plan, &d, g_host.stream, want_grad);
  if (rc == 0 && st.err == cudaSuccess && host_all_reduced) {
    void* ws = c->peer_workspace[c->rank];
    rc = hgf_fused_reduce_wait(ws, c->n_groups, c->dtype, c->world_size, c->epoch, g_host.stream);
    if (rc == 0)
      st.err = cudaMemcpyAsync(host_all_reduced, static_cast<char*>(ws) + hgf_fused_reduce_result_offset(c->n_groups, c->dtype, c->epoch),
                               static_cast<size_t>(c->n_groups) * w, cudaMemcpyDeviceToHost, g_host.stream);
  }
  st.finish();
  if (rc) return rc;
  if (st.err != cudaSuccess) return fail("host staging failed: %s", cudaGetErrorString(st.err));
  return 0;
}

extern "C" int hgf_filter_host(const hgf_plan_t* plan, const hgf_call_t* call, int32_t device) {
  return run_host(plan, call, device, false);
}
extern "C" int hgf_logp_grad_host(const hgf_plan_t* plan, const hgf_call_t* call, int32_t device) {
  return run_host(plan, call, device, true);
}
