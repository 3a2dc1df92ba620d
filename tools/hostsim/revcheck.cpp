// Host build of ONE generated reverse sweep (any topology / update scheme revgen.py covers), float64:
//   g++ -O2 -ffp-contract=off -fopenmp -shared -fPIC -I pyhgf_b200/csrc -I /usr/local/cuda/include -I <dir of the generated header>
//       -DREV_HEADER='<reverse_NAME.h>' -DREV_TYPE=Rev_NAME tools/hostsim/revcheck.cpp
// See tools/hostsim/revcheck.py (development tool + tests/test_device_equations_on_host.py).
#include "hgf_reverse.cuh"
#include REV_HEADER
#include <vector>

using namespace hgf;
using Rev = REV_TYPE<double>;

extern "C" int rev_n_slots() { return Rev::kSlots; }
extern "C" int rev_n_inputs() { return Rev::kInputs; }
// u [T][kInputs] shared by the instances, y [T] or null, attrs [kSlots], params [P][n]
extern "C" void rev_run(int n, int T, const double* u, const double* y, const double* time_steps, const double* attrs, int n_params, const int* slots,
                        const double* params, int response_fn, double max_posterior_precision, double* logp, double* grad) {
  Program prog{};
  prog.max_posterior_precision = max_posterior_precision;
  CallLists lists{};
  lists.n_params = n_params;
  lists.n_filter = n_params;
  for (int q = 0; q < n_params; ++q) { lists.param_slot[q] = static_cast<uint16_t>(slots[q]); lists.param_row[q] = static_cast<uint16_t>(q); }
  ScanArgs a{};
  a.n_instances = n; a.ld = n; a.n_steps = T;
  a.input_data = u; a.response_inputs = y; a.time_steps = time_steps; a.attributes = attrs; a.params = params;
  a.surprise_sum = logp; a.gradient = grad; a.out_scale = 1.0; a.response_fn = response_fn;
  std::vector<double> state(static_cast<size_t>(T) * Rev::kCarried * n);
#pragma omp parallel for schedule(dynamic, 16)
  for (int i = 0; i < n; ++i) {
    if (time_steps) reverse_instance<double, Rev, false>(prog, lists, a, i, true, state.data());
    else reverse_instance<double, Rev, true>(prog, lists, a, i, true, state.data());
  }
}
