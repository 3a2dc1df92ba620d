/* Plain-C caller of libpyhgf_cuda (no Python, no torch): the 3-level binary HGF of pyhgf's tutorial
 * (docs/source/notebooks/1.1-Binary_HGF.ipynb cell 20), N parameter samples, summed first-level binary surprise.
 *
 *   gcc -O2 -I include examples/c_abi_binary3.c -L pyhgf_b200/lib -lpyhgf_cuda -Wl,-rpath,$PWD/pyhgf_b200/lib -o c_abi_binary3
 *   ./c_abi_binary3 inputs.txt N out.txt        (inputs.txt: T lines of 0/1;  out.txt: N lines "omega_2 omega_3 surprise")
 *
 * The topology arrays below are what Network().add_nodes(kind="binary-state").add_nodes(value_children=0)
 * .add_nodes(volatility_children=1).get_network() holds (edges: typing.py:12-52; update sequence:
 * utils/get_update_sequence.py), written out by hand. */
#include <stdio.h>
#include <stdlib.h>

#include "pyhgf_cuda.h"

#define CHECK(call) do { if (call) { fprintf(stderr, "%s\n", hgf_last_error()); return 1; } } while (0)

int main(int argc, char** argv) {
  if (argc < 4) { fprintf(stderr, "usage: %s inputs.txt N out.txt\n", argv[0]); return 2; }
  FILE* f = fopen(argv[1], "r");
  if (!f) { perror(argv[1]); return 2; }
  float* u = malloc(sizeof(float) * 100000);
  int T = 0;
  while (T < 100000 && fscanf(f, "%f", &u[T]) == 1) ++T;
  fclose(f);
  const int N = atoi(argv[2]);

  /* node 0: binary state; node 1: its value parent; node 2: volatility parent of node 1 */
  const int32_t node_type[3] = {1, 2, 2};
  const int32_t vp_ptr[4] = {0, 1, 1, 1}, vp_idx[1] = {1};       /* value parents:        0 <- 1 */
  const int32_t volp_ptr[4] = {0, 0, 1, 1}, volp_idx[1] = {2};   /* volatility parents:   1 <- 2 */
  const int32_t vc_ptr[4] = {0, 0, 1, 1}, vc_idx[1] = {0}, vc_fn[1] = {HGF_CF_LINEAR};
  const int32_t volc_ptr[4] = {0, 0, 0, 1}, volc_idx[1] = {1};
  const int32_t pred_node[3] = {2, 1, 0}, pred_op[3] = {HGF_OP_PRED_CONT, HGF_OP_PRED_CONT, HGF_OP_PRED_BINARY};
  const int32_t upd_node[4] = {0, 1, 1, 2};
  const int32_t upd_op[4] = {HGF_OP_PE_BINARY, HGF_OP_POST_CONT, HGF_OP_PE_CONT, HGF_OP_POST_CONT_UNBOUNDED};
  const int32_t inputs[1] = {0};
  hgf_topology_t t = {0};
  t.n_nodes = 3; t.node_type = node_type;
  t.value_parents_ptr = vp_ptr; t.value_parents_idx = vp_idx;
  t.volatility_parents_ptr = volp_ptr; t.volatility_parents_idx = volp_idx;
  t.value_children_ptr = vc_ptr; t.value_children_idx = vc_idx; t.value_children_fn = vc_fn;
  t.volatility_children_ptr = volc_ptr; t.volatility_children_idx = volc_idx;
  t.n_prediction_steps = 3; t.prediction_node = pred_node; t.prediction_op = pred_op;
  t.n_update_steps = 4; t.update_node = upd_node; t.update_op = upd_op;
  t.n_inputs = 1; t.input_idxs = inputs; t.n_fed = 1;
  t.volatility_updates = 0; t.mean_field_updates = 0; t.max_posterior_precision = 1e10;
  hgf_plan_t* plan = NULL;
  CHECK(hgf_plan_create(&t, &plan));
  printf("kernel: %s, %d attribute slots\n", hgf_plan_kernel_name(plan), hgf_plan_n_slots(plan));

  /* initial attributes: node defaults of pyhgf/model/add_nodes.py:65-85, 330-340 (field order: hgf_program.h) */
  enum { MEAN = 0, EXPECTED_MEAN, PRECISION, EXPECTED_PRECISION, TONIC_VOLATILITY, TONIC_DRIFT, AUTOCONNECTION, OBSERVED };
  const int S = hgf_plan_n_slots(plan);
  float* attr = calloc(S, sizeof(float));
  attr[0 * HGF_N_FIELDS + EXPECTED_MEAN] = 0.5f; attr[0 * HGF_N_FIELDS + PRECISION] = 1.0f; attr[0 * HGF_N_FIELDS + EXPECTED_PRECISION] = 1.0f;
  attr[0 * HGF_N_FIELDS + OBSERVED] = 1.0f;
  for (int n = 1; n < 3; ++n) {
    float* a = attr + n * HGF_N_FIELDS;
    a[PRECISION] = 1.0f; a[EXPECTED_PRECISION] = 1.0f; a[TONIC_VOLATILITY] = -4.0f; a[AUTOCONNECTION] = 1.0f; a[OBSERVED] = 1.0f;
  }
  for (int kind = 0; kind < 4; ++kind)
    for (int n = 0; n < 3; ++n) {
      const int s = hgf_plan_slot_coupling(plan, kind, n, 0);
      if (s >= 0) attr[s] = 1.0f;  /* every coupling strength defaults to 1 */
    }
  attr[hgf_plan_slot_clipping(plan)] = 1e-6f;

  /* per-instance parameters: omega_2, omega_3 on a diagonal of [-6, -3] */
  float* params = malloc(sizeof(float) * 2 * N);
  for (int i = 0; i < N; ++i) {
    params[i] = -6.0f + 3.0f * i / (N > 1 ? N - 1 : 1);
    params[N + i] = -3.0f - 3.0f * i / (N > 1 ? N - 1 : 1);
  }
  const int32_t slots[2] = {1 * HGF_N_FIELDS + TONIC_VOLATILITY, 2 * HGF_N_FIELDS + TONIC_VOLATILITY};
  float* surprise = malloc(sizeof(float) * N);
  hgf_call_t c = {0};
  c.dtype = 32; c.n_instances = N; c.n_steps = T;
  c.input_data = u; c.attributes = attr;
  c.n_params = 2; c.param_slots = slots; c.params = params;
  c.response_fn = HGF_RESP_FIRST_LEVEL_BINARY;
  c.surprise_sum = surprise; c.out_scale = 1.0;
  CHECK(hgf_filter_host(plan, &c, 0));
  FILE* o = fopen(argv[3], "w");
  for (int i = 0; i < N; ++i) fprintf(o, "%.9g %.9g %.9g\n", params[i], params[N + i], surprise[i]);
  fclose(o);
  printf("launches: %lld\n", (long long)hgf_launch_count());
  hgf_plan_destroy(plan);
  return 0;
}
