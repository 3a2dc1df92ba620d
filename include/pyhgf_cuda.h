/* libpyhgf_cuda -- C ABI of the B200 (sm_100a) belief-propagation backend.
 *
 * The reference (ComputationalPsychiatry/pyhgf) has no FFI on this path: the
 * JAX implementation is traced Python (pyhgf/model/network.py:528-628 runs
 * `lax.scan(beliefs_propagation, ...)`), the Rust one is a PyO3 class
 * (src/model/network.rs:1415-1434).  These entry points are what a binding for
 * the path would call; INTEGRATION.md shows the ctypes and XLA-FFI stubs.
 *
 * Conventions: return 0 on success, non-zero on invalid arguments or a CUDA
 * error (text via hgf_last_error(), thread-local).  Numerical failure is NOT
 * an error: NaN / +inf flow in-band exactly as in the reference
 * (prediction/continuous.py:229-231, response.py:48).  The caller owns every
 * buffer.  Device entry points enqueue on `stream` and never synchronise; the
 * *_host entry points take host pointers, stage through the device and return
 * after the results are back in the caller's buffers.
 */
#ifndef PYHGF_CUDA_H_
#define PYHGF_CUDA_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HGF_ABI_VERSION 5

/* number of scalar attribute fields per node, and their order (slot = node * HGF_N_FIELDS + field).
 * Names are the reference's attribute keys, pyhgf/model/add_nodes.py:65-85, 159-184, 330-340. */
#define HGF_N_FIELDS 19
#define HGF_N_RESPONSE_PARAMETERS 4

typedef struct hgf_plan hgf_plan_t;

/* What pyhgf.Network.get_network() returns, flattened (pyhgf/model/network.py:769-776):
 * edges (pyhgf/typing/typing.py:12-52, CSR), the update sequence
 * (pyhgf/utils/get_update_sequence.py:51-351, opcodes below) and the constructor
 * constants (pyhgf/model/network.py:85-91). */
typedef struct {
  int32_t n_nodes;
  const int32_t* node_type;               /* [n_nodes] 1 binary, 2 continuous, 6 volatile */
  const int32_t* value_parents_ptr;       /* [n_nodes+1] */
  const int32_t* value_parents_idx;
  const int32_t* volatility_parents_ptr;  /* [n_nodes+1] */
  const int32_t* volatility_parents_idx;
  const int32_t* value_children_ptr;      /* [n_nodes+1] */
  const int32_t* value_children_idx;
  const int32_t* value_children_fn;       /* coupling function per value child: hgf_coupling_fn */
  const int32_t* volatility_children_ptr; /* [n_nodes+1] */
  const int32_t* volatility_children_idx;
  int32_t n_prediction_steps;
  const int32_t* prediction_node;
  const int32_t* prediction_op;           /* hgf_op */
  int32_t n_update_steps;
  const int32_t* update_node;
  const int32_t* update_op;
  int32_t n_inputs;                       /* Network.input_idxs, network.py:157-180 */
  const int32_t* input_idxs;
  int32_t n_fed;                          /* leading input nodes that receive a data column */
  int32_t volatility_updates;             /* 0 unbounded, 1 eHGF, 2 standard */
  int32_t mean_field_updates;
  double max_posterior_precision;
  /* Switching plans -- Network.input_custom_sequence (pyhgf/model/network.py:653-767: lax.switch over several update
   * sequences, one branch index per step).  n_branches <= 1: one schedule, as above.  Otherwise prediction_* /
   * update_* hold the branches' steps back to back (n_prediction_steps / n_update_steps are the totals) and branch b
   * owns steps [branch_prediction_ptr[b], branch_prediction_ptr[b+1]) and [branch_update_ptr[b], branch_update_ptr[b+1]).
   * At most HGF_MAX_BRANCHES; such a plan runs through the opcode interpreter and needs hgf_call_t::branch_idx. */
  int32_t n_branches;
  const int32_t* branch_prediction_ptr;   /* [n_branches + 1] */
  const int32_t* branch_update_ptr;       /* [n_branches + 1] */
} hgf_topology_t;

#define HGF_MAX_BRANCHES 8

enum hgf_coupling_fn { HGF_CF_LINEAR = 0, HGF_CF_IDENTITY = 1, HGF_CF_RELU = 2, HGF_CF_SIGMOID = 3, HGF_CF_TANH = 4,
                       HGF_CF_LEAKY_RELU = 5, HGF_CF_GELU = 6, HGF_CF_SIN = 7 };

enum hgf_op { HGF_OP_PRED_BINARY = 1, HGF_OP_PRED_CONT = 2, HGF_OP_PRED_CONT_MF = 3, HGF_OP_PRED_VOLATILE = 4,
              HGF_OP_PRED_VOLATILE_MF = 5, HGF_OP_PE_BINARY = 10, HGF_OP_PE_CONT = 11, HGF_OP_PE_VOLATILE = 12,
              HGF_OP_POST_CONT = 20, HGF_OP_POST_CONT_MF = 21, HGF_OP_POST_CONT_EHGF = 22, HGF_OP_POST_CONT_EHGF_MF = 23,
              HGF_OP_POST_CONT_UNBOUNDED = 24, HGF_OP_POST_VOLATILE = 25, HGF_OP_POST_VOLATILE_MF = 26 };

/* the built-in response functions of pyhgf/response.py:15-215 */
enum hgf_response { HGF_RESP_NONE = 0, HGF_RESP_FIRST_LEVEL_GAUSSIAN = 1, HGF_RESP_TOTAL_GAUSSIAN = 2,
                    HGF_RESP_FIRST_LEVEL_BINARY = 3, HGF_RESP_BINARY_SOFTMAX = 4,
                    HGF_RESP_BINARY_SOFTMAX_INVERSE_TEMPERATURE = 5 };

/* One batched evaluation: N independent instances of the plan's network, T time steps.
 * Replaces Network.input_data (+ Network.surprise / the vmap'd logp pattern) for the batch.
 * All arrays are dense; element type is float (dtype 32) or double (dtype 64). */
typedef struct {
  int32_t dtype;                 /* 32 or 64 */
  int64_t n_instances;           /* N */
  int32_t n_steps;               /* T */
  /* observations: instance n reads series n / input_group_size (0: one series shared by all) */
  int64_t input_group_size;
  const void* input_data;        /* [G][T][n_fed] */
  const uint8_t* observed;       /* [G][T][n_fed] or NULL (all observed) */
  const void* time_steps;        /* [T] or NULL (all 1) */
  /* attributes: the initial attribute vector shared by all instances, with per-instance overrides */
  const void* attributes;        /* [n_slots], see hgf_plan_n_slots / hgf_plan_slot_* */
  int32_t n_params;              /* P <= 16 */
  const int32_t* param_slots;    /* HOST [P]: slot each per-instance parameter overrides */
  const void* params;            /* [P][N] */
  /* response (surprise) function fused into the time loop */
  int32_t response_fn;           /* hgf_response */
  int64_t response_group_size;   /* grouping of response_inputs, like input_group_size */
  const void* response_inputs;   /* [G][T] or NULL */
  /* outputs (each may be NULL) */
  int32_t n_records;
  const int32_t* record_slots;   /* HOST [R]: attribute slots recorded at every step */
  void* trajectories;            /* [R][T][N] (time-major: stores coalesce over instances) */
  void* last_attributes;         /* [n_nodes*HGF_N_FIELDS][N] */
  void* surprise_sum;            /* [N]  out_scale * sum_t surprise  (out_scale = -1: log-probability) */
  void* surprise_pointwise;      /* [T][N] (unscaled) */
  void* gradient;                /* [P][N] d(surprise_sum)/d(params); hgf_logp_grad only */
  void* reduced;                 /* [n_groups] += surprise_sum of the instances of each group; caller zeroes it */
  const int32_t* group;          /* [N] group (e.g. chain) of each instance, or NULL: everything into reduced[0] */
  double out_scale;
  /* Fused cross-GPU reduction (optional; world_size <= 1: off).  One process per GPU; every rank allocates a
   * workspace of hgf_fused_reduce_bytes() in memory that its peers can map (CUDA VMM / torch symmetric memory),
   * zero-filled before the first launch, and passes all ranks' mappings.  The kernel then sums the per-group
   * (per-chain) results of ALL ranks itself: the last block of each rank adds its G partial sums into every
   * peer's workspace over NVLink (one multimem.red per value when a multicast mapping is given, else one red per
   * peer), publishes "launch `epoch` deposited" to every peer, and the kernel ENDS -- it does not wait for the
   * other ranks' sums, so a rank is never held back by its peers' launch skew, and no NCCL call follows the kernel.
   * The CONSUMER waits: hgf_fused_reduce_wait(workspace, ..., epoch, stream) enqueues a one-block kernel that
   * returns once every rank has deposited launch `epoch`; after it, workspace +
   * hgf_fused_reduce_result_offset(n_groups, dtype, epoch) holds the all-reduced [G] sums.  They stay valid for the
   * next HGF_FUSED_RING / 2 - 1 launches on this workspace; ranks may drift apart by as many launches (a launch
   * that is further ahead waits for the slowest peer before depositing).  `reduced` must be NULL; `epoch` counts
   * the launches on this workspace (0, 1, 2, ...) and must be the same on every rank. */
  int32_t n_groups;              /* G (also read by the *_host entry points: > 0 saves their scan of group[] for its maximum) */
  int32_t world_size;            /* W <= 8 */
  int32_t rank;
  uint32_t epoch;
  void* const* peer_workspace;   /* HOST [W]: device addresses, in THIS process, of every rank's workspace */
  void* multicast_workspace;     /* multicast address of the same workspace, or NULL */
  /* Generative mode (beliefs_propagation(observations="generative"), pyhgf/utils/beliefs_propagation.py:109-125,
   * pyhgf/utils/sample.py:18-116): instead of reading input_data, every input node observes a draw from its own
   * prediction -- Bernoulli(expected_mean) for binary nodes, Normal(expected_mean, 1/expected_precision) for
   * continuous ones (sample_node_distribution.py:34-40) -- with observed = 1.  As in the reference, all input nodes
   * of one step share one random draw.  The stream is Philox4x32-10 keyed by (seed, instance, step); instances are
   * the reference's n_predictions.  input_data / observed are ignored; hgf_filter only. */
  int32_t generative;
  uint64_t seed;
  /* Switching plans only: the branch taken at each step, [T] (shared by the instances; out-of-range values are clamped
   * like lax.switch).  Device pointer for hgf_filter / hgf_logp_grad, host pointer for the *_host entry points. */
  const int32_t* branch_idx;
  /* hgf_logp_grad only: how the gradient is computed.  0 (auto): forward-mode tangents carried in the filter's registers
   * (up to 4 filter parameters + 1 response parameter per pass); when that would take two or more passes -- or the plan's
   * kernels include no register-resident gradient kernel at all -- and the topology has a generated one, the checkpointed
   * REVERSE sweep (forward over time storing the carried beliefs, then backward with
   * the adjoint of every update: cost independent of the number of parameters).  1: forward-mode always.  2: the reverse
   * sweep (error if the plan's topology has none, or the call asks for trajectories / pointwise output / a fused
   * cross-GPU reduction, which only the forward-mode kernels produce). */
  int32_t gradient_mode;
} hgf_call_t;

int hgf_abi_version(void);
const char* hgf_last_error(void);

int hgf_plan_create(const hgf_topology_t* topology, hgf_plan_t** out);
void hgf_plan_destroy(hgf_plan_t* plan);

/* Register a register-resident specialisation built on demand for a user topology: `path` is a shared object
 * produced by pyhgf_b200/specialise.py (the generated topology accessor + the kernels instantiated for it,
 * compiled for sm_100a with this library's own headers).  Plans created afterwards dispatch to it when their
 * network answers every topology query the way the specialisation was generated; everything else keeps running
 * through the opcode interpreter.  The object stays loaded for the life of the process. */
int hgf_register_specialisation(const char* path);

/* attribute-vector layout */
int32_t hgf_plan_n_slots(const hgf_plan_t* plan);
/* kind: 0 value_coupling_parents, 1 volatility_coupling_parents, 2 value_coupling_children,
 * 3 volatility_coupling_children */
int32_t hgf_plan_slot_coupling(const hgf_plan_t* plan, int32_t kind, int32_t node, int32_t position);
int32_t hgf_plan_slot_clipping(const hgf_plan_t* plan);   /* attributes[-1]["precision_clipping_value"] */
int32_t hgf_plan_slot_response(const hgf_plan_t* plan);   /* first response-function parameter */
/* name of the kernel variant the plan dispatches to ("interpreter" or a register-resident specialisation) */
const char* hgf_plan_kernel_name(const hgf_plan_t* plan);

/* Forward filter (+ fused surprise).  Device pointers; enqueued on `stream` (a cudaStream_t). */
int hgf_filter(const hgf_plan_t* plan, const hgf_call_t* call, void* stream);
/* Forward filter + d(surprise_sum)/d(params) in the same pass.  Device pointers. */
int hgf_logp_grad(const hgf_plan_t* plan, const hgf_call_t* call, void* stream);

/* hgf_logp_grad's reverse sweep (gradient_mode = 2) for callers that cannot allocate: code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc.  (No counterpart in the reference.)
 * hgf_logp_grad takes the sweep's checkpoint array from cudaMallocAsync; this entry point keeps it in
 * a static device array of HGF_NOALLOC_SCRATCH_BYTES (32 MiB, held by every GPU that loads the library until the
 * process exits) and calls no cudaMalloc, cudaFree or allocating library routine.  Scope: the 3-level binary HGF
 * (hgf_plan_kernel_name() == "static:binary3"), float32, a per-trial response function, outputs surprise_sum /
 * gradient / reduced; n_steps * n_instances <= HGF_NOALLOC_MAX_STEP_INSTANCES (e.g. 320 trials x 6 553 instances).
 * Anything else is rejected with an error (hgf_last_error()).  Same inputs, outputs and results as hgf_logp_grad.
 * The scratch is shared by every call on a device and keeps the last call's contents; nothing in it needs clearing
 * (a launch reads back only checkpoints it has written itself), but two calls that OVERLAP on the GPU (different
 * streams) race on it: serialise them.  The first call on each device looks up the array's device address
 * (cudaGetSymbolAddress); make that call outside stream capture. */
#define HGF_NOALLOC_SCRATCH_BYTES 33554432
#define HGF_NOALLOC_MAX_STEP_INSTANCES 2097152
int hgf_logp_grad_reverse_noalloc(const hgf_plan_t* plan, const hgf_call_t* call, void* stream);

/* hgf_filter / hgf_logp_grad with every data pointer in HOST memory (pageable or pinned); copies in, runs, copies out, synchronises.
 * With a fused cross-GPU reduction (world_size > 1; peer_workspace stays a table of DEVICE addresses) a non-NULL host
 * `reduced` receives the sums over ALL ranks: the call waits for every rank's deposit before copying them out. */
int hgf_filter_host(const hgf_plan_t* plan, const hgf_call_t* call, int32_t device);
int hgf_logp_grad_host(const hgf_plan_t* plan, const hgf_call_t* call, int32_t device);

/* fused cross-GPU reduction workspace: size, where the all-reduced [n_groups] sums of a launch are, and the
 * consumer's wait for them (enqueued on `stream`; see hgf_call_t::peer_workspace) */
#define HGF_FUSED_RING 16
int64_t hgf_fused_reduce_bytes(int32_t n_groups, int32_t dtype);
int64_t hgf_fused_reduce_result_offset(int32_t n_groups, int32_t dtype, uint32_t epoch);
int hgf_fused_reduce_wait(void* workspace, int32_t n_groups, int32_t dtype, int32_t world_size, uint32_t epoch, void* stream);

/* Exponential-family state node with a fixed learning rate ("generalised Bayesian filtering":
 * add_nodes(kind="ef-state", learning="generalised-filtering"), pyhgf/model/add_nodes.py:360-470;
 * prediction_error_update_exponential_family_fixed, pyhgf/updates/prediction_error/exponential.py:13-60).
 * The node has no parents: at every step   xis <- xis + (t(x) - xis) / (1 + nu)   where observed, with the
 * sufficient statistics t(x) = [x, vech_rowwise_lower(x x^T)] of a (multivariate) normal (pyhgf/math.py:90-92,
 * MultivariateNormal.sufficient_statistics_from_observations), K = d + d (d + 1) / 2 values.  N instances differ
 * in their learning rate nu (a sweep) and/or their series. */
typedef struct {
  int32_t dtype;               /* 32 or 64 */
  int64_t n_instances;         /* N */
  int32_t n_steps;             /* T */
  int32_t dimension;           /* d <= 8 */
  int64_t input_group_size;    /* instance n reads series n / input_group_size (0: one shared series) */
  const void* input_data;      /* [G][T][d] */
  const uint8_t* observed;     /* [G][T] or NULL */
  const void* nus;             /* [N] */
  const void* xis0;            /* [K] initial sufficient statistics, shared */
  void* xis_trajectory;        /* [K][T][N] or NULL */
  void* xis_last;              /* [K][N] or NULL */
  /* learning = "hgf-1|2|3" (add_nodes.py:438-452, posterior_update_exponential_family_dynamic,
   * pyhgf/updates/posterior/exponential.py): the learning rate of step t is implied by the continuous HGFs that track the
   * sufficient statistics, nu_t = mean_k (A_k.mean - A_k.expected_mean) / (B_k.mean - B_k.expected_mean), A_k the node
   * fed with statistic k and B_k its value parent.  implied = their trajectories as hgf_filter records them,
   * [4][K][T][N] in the order A.mean, A.expected_mean, B.mean, B.expected_mean; nus is then ignored. */
  const void* implied;         /* NULL: fixed learning rate (nus) */
  void* nus_trajectory;        /* [T][N] or NULL: the implied nu_t */
} hgf_ef_call_t;
int hgf_ef_filter(const hgf_ef_call_t* call, void* stream);       /* device pointers */

/* Categorical state node (pyhgf/model/add_nodes.py:473-545, pyhgf/utils/fill_categorical_state_node.py:13-81): one binary
 * HGF per category under the node, run by hgf_filter like any other nodes (the binary nodes are the plan's inputs, one
 * observation column per category).  This call is the node's own step, categorical_state_update
 * (pyhgf/updates/posterior/categorical.py:12-91), for every step and instance at once, from the recorded trajectories
 * of the implied nodes: alpha_t = nu_t * mu_hat_t + 1 with nu_t = (x_t - mu_hat_t) / (sigmoid(mu_parent_t) - mu_hat_t) - 1,
 * kl_t = KL(Dir(alpha_{t-1}) || Dir(alpha_t)) (pyhgf/math.py:307-341), surprise_t = sum_k binary_surprise(x_tk, mu_hat_tk).
 * All pointers are device pointers; trajectories are laid out as hgf_filter writes them, [category][T][N]. */
#define HGF_MAX_CATEGORIES 16
typedef struct {
  int32_t dtype;               /* 32 or 64 */
  int64_t n_instances;         /* N */
  int32_t n_steps;             /* T */
  int32_t n_categories;        /* n, 2..HGF_MAX_CATEGORIES */
  int64_t input_group_size;    /* 0: one observation series for all instances; g > 0: instance i reads series i / g */
  const void* observations;    /* [G][T][n] the observed category, one-hot */
  const void* expected_mean;   /* [n][T][N] expected_mean of the binary nodes */
  const void* parent_mean;     /* [n][T][N] posterior mean of their continuous value parents */
  const void* alpha0;          /* [n] the concentration before the first step (attributes[node]["alpha"]) */
  void* alpha;                 /* [n][T][N] or NULL */
  void* mean;                  /* [n][T][N] or NULL: sigmoid(parent_mean), the node's "mean" after the update */
  void* kl_divergence;         /* [T][N] */
  void* surprise;              /* [T][N] */
} hgf_categorical_call_t;

int hgf_categorical_update(const hgf_categorical_call_t* call, void* stream);

/* number of kernels this library has launched in the calling process (bench.py's gpu_launches) */
int64_t hgf_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif  /* PYHGF_CUDA_H_ */
