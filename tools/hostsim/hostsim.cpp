// Numerics test-bed (development tool, not part of the library): the DEVICE equations of pyhgf_b200/csrc
// (hgf_math.cuh, hgf_ops.cuh, hgf_scan.cuh -- all __host__ __device__) compiled for the HOST, so that a math
// policy can be studied without a GPU.  With -DHGF_HOST_EMULATE_SFU the SFU primitives are emulated inside their
// documented error bounds (hgf_math.cuh) and the policy knobs (-DHGF_DIV_MODE=..., -DHGF_RESP_MODE=...,
// -DHGF_EXACT_DIV_SITES=...) act as they do on the device; without it every primitive is IEEE / libm.
//   g++ -O2 -ffp-contract=off -shared -fPIC -I pyhgf_b200/csrc -I /usr/local/cuda/include [-D...] tools/hostsim/hostsim.cpp
// See tools/hostsim/run.py.
#include "hgf_scan.cuh"
#include <generated/topo_binary3.h>
#include <generated/topo_continuous3.h>
#include "hgf_reverse.cuh"
#include <generated/reverse_binary3.h>
#include <generated/reverse_continuous3.h>
#ifdef HOSTSIM_GHGF16
#include <generated/reverse_ghgf16.h>  // 12 000 generated lines: only the test that needs it pays the compile (-DHOSTSIM_GHGF16)
#endif
#include <vector>

using namespace hgf;

template <class Topo, int PF, int PR>
static void run(int n_inst, int n_steps, int group, const float* u, const float* y, const float* attrs, int n_params, const int* slots,
                const float* params, int response_fn, float* logp, float* grad) {
  Program prog{};
  prog.max_posterior_precision = 1e10;
  CallLists lists{};
  lists.n_params = n_params;
  lists.n_filter = PR ? n_params - 1 : n_params;
  for (int q = 0; q < n_params; ++q) { lists.param_slot[q] = static_cast<uint16_t>(slots[q]); lists.param_row[q] = static_cast<uint16_t>(q); }
  ScanArgs a{};
  a.n_instances = n_inst; a.ld = n_inst; a.n_steps = n_steps;
  a.input_group_size = group; a.response_group_size = group;
  a.input_data = u; a.response_inputs = y; a.attributes = attrs; a.params = params;
  a.surprise_sum = logp; a.gradient = (PF + PR) ? grad : nullptr;
  a.out_scale = 1.0; a.response_fn = response_fn; a.tangent_offset = 0; a.resp_index = PR ? n_params - 1 : -1;
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = 0; i < n_inst; ++i) scan_instance<Dual<float, PF>, Topo, Topo::kSlots, PR, true, false>(prog, lists, a, i, true);
}

extern "C" void sim_binary3_grad(int n, int T, int group, const float* u, const float* y, const float* attrs, const int* slots,
                                 const float* params, float* logp, float* grad) {
  run<Topo_binary3, 2, 1>(n, T, group, u, y, attrs, 3, slots, params, RESP_BINARY_SOFTMAX_INVERSE_TEMPERATURE, logp, grad);
}
// the ConstAttributes<> flavour of the same kernel (no structural attribute is differentiated)
extern "C" void sim_binary3_grad_const(int n, int T, int group, const float* u, const float* y, const float* attrs, const int* slots,
                                       const float* params, float* logp, float* grad) {
  run<ConstAttributes<Topo_binary3>, 2, 1>(n, T, group, u, y, attrs, 3, slots, params, RESP_BINARY_SOFTMAX_INVERSE_TEMPERATURE, logp, grad);
}
// the generated reverse sweep (hgf_reverse.cuh) of the same evaluation; params are ALL per-instance parameters [P][n]
template <class F, class Rev>
static void run_reverse(int n_inst, int n_steps, int group, const F* u, const F* y, const F* attrs, int n_params, const int* slots, const F* params,
                        int response_fn, F* logp, F* grad) {
  Program prog{};
  prog.max_posterior_precision = 1e10;
  CallLists lists{};
  lists.n_params = n_params;
  lists.n_filter = n_params;
  for (int q = 0; q < n_params; ++q) { lists.param_slot[q] = static_cast<uint16_t>(slots[q]); lists.param_row[q] = static_cast<uint16_t>(q); }
  ScanArgs a{};
  a.n_instances = n_inst; a.ld = n_inst; a.n_steps = n_steps;
  a.input_group_size = group; a.response_group_size = group;
  a.input_data = u; a.response_inputs = y; a.attributes = attrs; a.params = params;
  a.surprise_sum = logp; a.gradient = grad; a.out_scale = 1.0; a.response_fn = response_fn;
  std::vector<F> state(static_cast<size_t>(n_steps) * Rev::kCarried * n_inst);
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = 0; i < n_inst; ++i) reverse_instance<F, Rev, true>(prog, lists, a, i, true, state.data());
}
extern "C" void sim_binary3_reverse(int n, int T, int group, const float* u, const float* y, const float* attrs, int n_params, const int* slots,
                                    const float* params, int response_fn, float* logp, float* grad) {
  run_reverse<float, Rev_binary3<float>>(n, T, group, u, y, attrs, n_params, slots, params, response_fn, logp, grad);
}
extern "C" void sim_binary3_reverse_f64(int n, int T, int group, const double* u, const double* y, const double* attrs, int n_params, const int* slots,
                                        const double* params, int response_fn, double* logp, double* grad) {
  run_reverse<double, Rev_binary3<double>>(n, T, group, u, y, attrs, n_params, slots, params, response_fn, logp, grad);
}
extern "C" void sim_continuous3_reverse_f64(int n, int T, const double* u, const double* attrs, int n_params, const int* slots, const double* params,
                                            double* logp, double* grad) {
  run_reverse<double, Rev_continuous3<double>>(n, T, 0, u, nullptr, attrs, n_params, slots, params, RESP_FIRST_LEVEL_GAUSSIAN, logp, grad);
}
#ifdef HOSTSIM_GHGF16
// BASELINE config 5's 16-node network (continuous + volatile-state nodes), u [T][4] shared by the instances
extern "C" void sim_ghgf16_reverse_f64(int n, int T, const double* u, const double* attrs, int n_params, const int* slots, const double* params,
                                       int response_fn, double* logp, double* grad) {
  run_reverse<double, Rev_ghgf16<double>>(n, T, 0, u, nullptr, attrs, n_params, slots, params, response_fn, logp, grad);
}
#endif
extern "C" void sim_binary3(int n, int T, int group, const float* u, const float* y, const float* attrs, const int* slots,
                            const float* params, int response_fn, float* logp) {
  run<Topo_binary3, 0, 0>(n, T, group, u, y, attrs, 3, slots, params, response_fn, logp, nullptr);
}
extern "C" void sim_continuous3(int n, int T, const float* u, const float* attrs, const int* slots, const float* params, float* logp) {
  run<Topo_continuous3, 0, 0>(n, T, 0, u, nullptr, attrs, 3, slots, params, RESP_FIRST_LEVEL_GAUSSIAN, logp, nullptr);
}
